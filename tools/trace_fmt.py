"""Format the last [gemm trace] block of a log as per-role tables (cycles since CTA start)."""
import sys, collections
lines = [l.split() for l in open(sys.argv[1]) if l.startswith("[gemm trace]")]
# keep the last repetition: split where role index drops back to 0 after being >0
blocks, cur, last_r = [], [], -1
for l in lines:
    r = int(l[2])
    if r < last_r: blocks.append(cur); cur = []
    cur.append(l); last_r = r
blocks.append(cur)
ev = collections.defaultdict(dict)
for l in blocks[-1]:
    r, i, e, t = int(l[2]), int(l[3]), int(l[4]), int(l[5])
    ev[(r, i)][e] = t
names = {**{o: "issuer%d [accempty actfull afull0 issued0 afull1 issued1]" % o for o in range(4)},
         **{4 + o: "dequant%d [wfull lds aempty0 afull0> aempty1 afull1>]" % o for o in range(4)},
         8: "epilogue [start accfull ld_done fma_done]", 9: "producer_w [issue]", 10: "producer_a [issue]"}
for r in range(11):
    print(names[r])
    for i in range(64):
        if (r, i) in ev:
            d = ev[(r, i)]
            print("  %2d: " % i + " ".join("%6d" % d[e] if e in d else "     -" for e in range(max(d) + 1)))
