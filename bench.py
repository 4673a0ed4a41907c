#!/usr/bin/env python
"""Headline benchmark: decode tokens/s of Llama-3-8B W4A16 g128 (gptq_marlin layout), random-init, synthetic step inputs.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--batch B] [--ctx L] [--model ...]

A "step" is ONE decode step of the whole model for `batch` sequences whose KV cache holds `ctx` tokens: 32 x
(rmsnorm, qkv GEMM, rope, reshape_and_cache, paged attention, o GEMM, add+rmsnorm, gate_up GEMM, silu*mul, down GEMM),
final norm, bf16 lm_head, greedy argmax -- BASELINE.json configs[1].  Geometry (ctx 256 = 128-token prompt + 128 decoded)
is configs[0]'s, batch 64 the top of "bs 1-64".  N > 1 (torchrun) = tensor parallel TP=N of the same model (strong scaling).

One JSON line on stdout (rank 0).  `value` = device-timed graph replays with inputs resident in HBM; `e2e` = the same
step through DecodeEngine.step() with pinned HOST inputs copied in and the sampled ids copied out every step;
`roofline` = the Marlin W4A16 GEMM (dominant kernel) timed alone over all 128 per-step launches with CUDA events;
`cpu_baseline` = the host-core restatement (oracle/cpu_decode.py) on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=20)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--batch", type=int, default=64)
    p.add_argument("--ctx", type=int, default=256)
    p.add_argument("--model", default="llama-3-8b", choices=["llama-3-8b", "llama-3-70b", "tiny"])
    p.add_argument("--quant", default="w4a16", choices=["w4a16", "w4a16_24", "fp8", "fp8_static", "int8"])
    p.add_argument("--kv", default="auto", choices=["auto", "fp8", "fp8_e5m2"])
    p.add_argument("--layers", type=int, default=None, help="debug: truncate the model (makes the number INVALID)")
    p.add_argument("--allreduce", default="auto", choices=["auto", "custom", "nccl"],
                   help="N>1: row-parallel all-reduce through the one-shot NVLink kernel (custom), NCCL, or custom with "
                        "NCCL as the loud fallback when peer memory cannot be mapped (auto)")
    p.add_argument("--no-fuse", action="store_true", help="N>1: keep GEMM and all-reduce as two kernels")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-sweep", action="store_true",
                   help="skip the batch sweep (tok/s at batch 1,2,4,8,16,32 + GEMM roofline fraction at batch 1 and 16)")
    p.add_argument("--no-extra-models", action="store_true",
                   help="N>1: skip the Llama-3-70B W4A16 / FP8 + fp8-KV ms-per-step lines at this TP degree")
    p.add_argument("--profile-step", action="store_true",
                   help="bracket ONE extra decode step with cudaProfilerStart/Stop (use with ncu --profile-from-start off)")
    return p.parse_args()


def shape_of(name):
    from nm_vllm_b200 import decoder
    return {"llama-3-8b": decoder.LLAMA3_8B, "llama-3-70b": decoder.LLAMA3_70B, "tiny": decoder.TINY}[name]


def workload_name(a):
    return "%s %s g128 decode, batch %d, ctx %d, kv %s" % (a.model, a.quant, a.batch, a.ctx, a.kv)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.p = [], None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                      stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.p is None:
            return None
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=2)
        except Exception:
            self.p.kill()
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
            except Exception:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return None
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def run_reference(a):
    """Reference arm: the path on the box's host cores (see oracle/cpu_decode.py), rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import torch
    from oracle import cpu_decode
    shape = shape_of(a.model)
    ts, last = [], None
    for i in range(a.warmup + a.steps):
        last = cpu_decode.time_decode(shape, a.batch, a.ctx, layer_reps=1, threads=os.cpu_count())
        if i >= a.warmup:
            ts.append(last["step_s"])
        if sum(ts) > 150:  # keep the whole run within minutes
            break
    step_s = sum(ts) / len(ts)
    v = a.batch / step_s
    out = {"impl": "reference", "metric": "decode_tokens_per_s", "value": v, "unit": "tok/s", "n_gpus": a.gpus,
           "steps": len(ts), "warmup": a.warmup, "ms_per_step": step_s * 1e3, "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
           "config": {"workload": workload_name(a), "parallelism": "host cores"},
           "cpu_baseline": {"value": v, "unit": "tok/s", "cores": last["cores"], "host_cpu_count": os.cpu_count(),
                            "kind": last["kind"], "sample": last["sample"]},
           "e2e": {"value": v, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out), flush=True)


_OUR_KERNELS_PER_CALL = {  # CUDA launches of OUR kernels behind one operator call
    "rms_norm": 1, "fused_add_rms_norm": 1, "silu_and_mul": 1, "rotary_embedding": 1, "reshape_and_cache": 1,
    "paged_attention_v1": 1, "paged_attention_v2": 2, "gptq_marlin_gemm": 1, "gptq_marlin_gemm_silu_and_mul": 1, "gptq_marlin_24_gemm": 1,
    "cutlass_scaled_mm": 1, "scaled_fp8_quant": 2, "scaled_int8_quant": 1, "greedy_sample": 1,
    "rope_cache_paged_attention": 1,   # + the v2 merge kernel when exp_sums is given (counted in the wrapper)
    "rms_norm_quant": 1, "silu_and_mul_quant": 1,   # (also with kind "fp8_dynamic": the grid-wide maximum is inside the launch)
}


def count_launches(fn):
    """Runs fn() once eagerly with counting wrappers around the operator module; returns our kernel launches."""
    from nm_vllm_b200 import _custom_ops as ops
    counts, saved = {}, {}
    for name, per in _OUR_KERNELS_PER_CALL.items():
        orig = getattr(ops, name)
        saved[name] = orig

        def wrap(*args, __o=orig, __n=name, __p=per, **kw):
            r = __o(*args, **kw)
            if __n == "rope_cache_paged_attention":
                if r:   # False = nothing launched (shape outside the fast path)
                    counts[__n] = counts.get(__n, 0) + (2 if args[1] is not None else 1)
            elif __n == "scaled_fp8_quant":   # dynamic: absmax + convert; static scale given: one kernel
                static = (len(args) > 1 and args[1] is not None) or kw.get("scale") is not None
                counts[__n] = counts.get(__n, 0) + (1 if static else 2)
            else:
                counts[__n] = counts.get(__n, 0) + __p
            return r
        setattr(ops, name, wrap)
    try:
        fn()
    finally:
        for name, orig in saved.items():
            setattr(ops, name, orig)
    return sum(counts.values()), counts


def gemm_roofline(model, batch, peaks, brief=False):
    """The dominant kernel timed ALONE: every Marlin GEMM of one step (4 per layer, distinct weights -> cold L2: the
    packed weights of one pass are 3.6 GB vs 126 MB of L2), CUDA events on the launching stream, 3 passes, median."""
    import torch
    dev = model.device
    lin = [(L[k]) for L in model.layers for k in ("qkv", "o", "gate_up", "down")]
    xs = {}
    for q in lin:
        if q.K not in xs:
            xs[q.K] = torch.randn((batch, q.K), device=dev, dtype=model.dtype)
    alg = sum(q.weight_bytes() + 2 * batch * (q.K + q.N) for q in lin)
    for q in lin[:8]:
        q(xs[q.K])
    torch.cuda.synchronize()
    times = []
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for q in lin:
            q(xs[q.K])
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1) * 1e-3)
    times.sort()
    t = times[1]
    # the same pass replayed from a CUDA graph removes host launch gaps from the per-launch duration
    st = torch.cuda.Stream(device=dev)
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(st):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            for q in lin:
                q(xs[q.K])
        g.replay()
        torch.cuda.synchronize()
        gt = []
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            g.replay()
            e1.record(st)
            torch.cuda.synchronize()
            gt.append(e0.elapsed_time(e1) * 1e-3)
    gt.sort()
    t = min(t, gt[len(gt) // 2])
    peak = peaks.get("hbm_gbs", 6650.0)
    ach = alg / t / 1e9
    if brief:
        return {"achieved": ach, "frac": ach / peak, "avg_launch_us": t / len(lin) * 1e6, "gemm_ms_per_step": t * 1e3}
    kernel = {"w4a16": "marlin_gemm_kernel (gptq_marlin_gemm W4A16 g128)",
              "w4a16_24": "marlin24_gemm_kernel (gptq_marlin_24_gemm 2:4 W4A16 g128)",
              "fp8": "scaled_mm_kernel (cutlass_scaled_mm fp8, per-token x per-channel)",
              "fp8_static": "scaled_mm_kernel (cutlass_scaled_mm fp8, per-tensor x per-channel)",
              "int8": "scaled_mm_kernel (cutlass_scaled_mm int8, per-token x per-channel)"}.get(lin[0].kind, lin[0].kind)
    return {"bound": "hbm", "kernel": kernel, "achieved": ach, "peak": peak,
            "unit": "GB/s", "frac": ach / peak,
            # dram bytes per launch are NOT measured in this run (ncu only): see profiles/r1_gemm_traffic.json for the
            # N=1 batch-64 capture (29.3 MB per launch vs 30.3 MB algorithmic)
            "traffic": None,
            "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650",
            "launches": len(lin), "alg_bytes_per_launch": alg / len(lin), "avg_launch_us": t / len(lin) * 1e6,
            "gemm_ms_per_step": t * 1e3}


def tp_parity_check(model, eng, batch, dev, rank, world):
    """N > 1 self-check BEFORE timing (the driver's 1-GPU test box cannot run the multi-rank kernels): the NVLink
    all-reduce kernel and the fused GEMM + all-reduce against an fp32 sum of the same per-rank inputs through NCCL
    (tests/distributed/test_custom_all_reduce.py:29-131 does the same with known tensors), and the sampled ids of one
    decode step identical on every rank.  Any mismatch makes the run exit non-zero."""
    import torch
    import torch.distributed as dist
    res = {"allreduce_kernel": "not used", "fused_gemm_allreduce": "not used"}
    ok = True
    g = torch.Generator(device=dev).manual_seed(1000 + rank)

    def close(y, ref32):
        # both sides sum bf16 inputs in fp32 (different order) and round once: at most one ulp of the output dtype apart
        tol = ref32.abs() * (2.0 ** -7 if y.dtype == torch.bfloat16 else 2.0 ** -10) + 1e-5
        bad = ((y.float() - ref32).abs() > tol).sum()
        return int(bad) == 0

    hidden = model.shape.hidden
    x = torch.randn((batch, hidden), generator=g, device=dev).to(model.dtype)
    ref = x.float().clone()
    dist.all_reduce(ref)
    if model.custom_ar is not None and model.custom_ar.should_custom_ar(x):
        y = model.custom_ar.all_reduce(x)
        same = [torch.empty_like(y) for _ in range(world)]
        dist.all_gather(same, y)
        identical = all(torch.equal(same[0], t) for t in same)
        good = close(y, ref) and identical
        res["allreduce_kernel"] = "ok" if good else "MISMATCH"
        ok = ok and good
    lin = model.layers[0]["o"]
    a = torch.randn((batch, lin.K), generator=g, device=dev).to(model.dtype)
    part = lin(a)
    ref2 = part.float().clone()
    dist.all_reduce(ref2)
    if (model.custom_ar is not None and model.fuse_gemm_all_reduce and lin.kind == "w4a16"
            and model.custom_ar.fuse_next_gemm(batch, lin.N)):
        y2 = lin(a)
        good = close(y2, ref2)
        res["fused_gemm_allreduce"] = "ok" if good else "MISMATCH"
        ok = ok and good
    eng.run_device()
    torch.cuda.synchronize()
    ids = [torch.empty_like(eng.d_out) for _ in range(world)]
    dist.all_gather(ids, eng.d_out)
    same_ids = all(torch.equal(ids[0], t) for t in ids)
    res["sampled_ids_identical_on_all_ranks"] = bool(same_ids)
    ok = ok and same_ids
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    res["ok"] = bool(int(flag))
    return res


def time_extra_model(a, decoder, shape, quant, kv, dev, world, rank, use_custom_ar, steps):
    """ms per decode step (device-timed CUDA-graph replays, max over ranks) of another model at this TP degree."""
    import torch
    import torch.distributed as dist
    dtype = torch.float16 if quant == "w4a16_24" else torch.bfloat16
    nb_seq = (a.ctx + 15) // 16
    m = decoder.SyntheticLlama(shape, quant=quant, dtype=dtype, device=dev, tp_size=world, tp_rank=rank,
                               kv_cache_dtype=kv, kv_scale=1.0 if kv == "auto" else 0.0376,
                               num_blocks=max(a.batch * nb_seq * 2, 1024), max_pos=max(a.ctx + 1, 512))
    if use_custom_ar is not None:
        m.custom_ar = use_custom_ar
    m.fuse_gemm_all_reduce = not a.no_fuse
    eng = decoder.DecodeEngine(m, a.batch, a.ctx)
    eng.load_inputs(eng.host_inputs(a.ctx, seed=1))
    eng.prepare(a.ctx)
    for _ in range(3):
        eng.run_device()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        eng.run_device()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / steps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    floor = (m.linear_weight_bytes() + m.lm_head.weight_bytes() + m.kv_read_bytes([a.ctx] * a.batch))
    return {"ms_per_step": float(t), "tok_s": a.batch / float(t) * 1e3, "hbm_floor_bytes_per_rank": floor, "layers": m.nlayers}


def run_b200(a):
    import torch
    import torch.distributed as dist
    from nm_vllm_b200 import decoder, load_extension
    load_extension()  # fails loudly when the sm_100a extension is missing: there is no fallback
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --impl b200 needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    dev = "cuda:%d" % local
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(dev))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    shape = shape_of(a.model)
    dtype = torch.float16 if a.quant == "w4a16_24" else torch.bfloat16
    nb_seq = (a.ctx + 15) // 16
    num_blocks = max(a.batch * nb_seq * 2, 1024)
    model = decoder.SyntheticLlama(shape, quant=a.quant, dtype=dtype, device=dev, tp_size=world, tp_rank=rank,
                                   kv_cache_dtype=a.kv, kv_scale=1.0 if a.kv == "auto" else 0.0376,
                                   num_blocks=num_blocks, max_pos=max(a.ctx + 1, 2048 + 16), layers=a.layers)
    ar_kind = "none"
    model.fuse_gemm_all_reduce = not a.no_fuse
    if world > 1:
        ar_kind = "nccl"
        if a.allreduce != "nccl":
            try:
                model.enable_custom_all_reduce()
                ar_kind = "one-shot NVLink peer-memory kernel (csrc/all_reduce.cu), NCCL for logits all-gather"
            except Exception as e:  # peer memory not mappable on this box
                if a.allreduce == "custom":
                    raise
                sys.stderr.write("[bench] rank %d: custom all-reduce unavailable (%s); using NCCL\n" % (rank, e))
            # all ranks must agree, otherwise the collectives would not match
            ok = torch.tensor([1 if model.custom_ar is not None else 0], device=dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if int(ok) == 0:
                model.custom_ar = None
                ar_kind = "nccl"
    eng = decoder.DecodeEngine(model, a.batch, a.ctx)
    hi = eng.host_inputs(a.ctx, seed=1)
    eng.load_inputs(hi)
    torch.cuda.synchronize()
    model.n_custom_ar_calls = 0
    launches, per_op = count_launches(lambda: eng._device_step(a.ctx))
    if model.custom_ar is not None:   # stand-alone all-reduce kernels (0 when every row-parallel GEMM fused its own)
        per_op["custom_all_reduce"] = model.n_custom_ar_calls
        launches += model.n_custom_ar_calls
        if model.n_custom_ar_calls == 0:
            ar_kind = "fused into the row-parallel W4A16 GEMM epilogue (LL exchange over NVLink peer memory)"
    eng.prepare(a.ctx)
    parity = None
    if world > 1:
        parity = tp_parity_check(model, eng, a.batch, dev, rank, world)
        if not parity["ok"]:
            if rank == 0:
                print(json.dumps({"parity_check": parity, "error": "multi-GPU parity check failed; nothing timed"}), flush=True)
            sys.stdout.flush()
            os._exit(3)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing --------------------------------------------------------------------------------
    for _ in range(max(a.warmup, 3)):
        eng.run_device()
    barrier()
    clocks = ClockSampler(local) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        eng.run_device()
    e1.record()
    barrier()
    t_dev = e0.elapsed_time(e1) * 1e-3
    if a.profile_step:
        torch.cuda.synchronize()
        torch.cuda.cudart().cudaProfilerStart()
        eng.run_device()
        torch.cuda.synchronize()
        torch.cuda.cudart().cudaProfilerStop()
    # ---- end to end: pinned host inputs in, sampled ids out, every step -------------------------------------------
    his = [eng.host_inputs(a.ctx, seed=10 + i) for i in range(4)]
    for i in range(3):
        eng.step(his[i % 4])
    barrier()
    e0.record()
    for i in range(a.steps):
        eng.step(his[i % 4])
    e1.record()
    barrier()
    t_e2e = e0.elapsed_time(e1) * 1e-3
    clk = clocks.stop() if clocks else None
    if world > 1:
        tt = torch.tensor([t_dev, t_e2e], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t_dev, t_e2e = float(tt[0]), float(tt[1])
    tok = a.batch * a.steps
    out = {"metric": "decode_tokens_per_s", "value": tok / t_dev, "unit": "tok/s", "n_gpus": world, "steps": a.steps,
           "warmup": max(a.warmup, 3), "ms_per_step": t_dev / a.steps * 1e3, "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "bf16" if dtype == torch.bfloat16 else "f16",
           "data": "synthetic",
           "config": {"workload": workload_name(a), "parallelism": "tp%d" % world, "layers": model.nlayers,
                      "allreduce": ar_kind,
                      "l2": "inputs larger than L2: %.2f GB of packed weights + lm_head and %.2f GB of KV are streamed "
                            "every step (L2 = 126 MB), no flush needed" %
                            ((model.linear_weight_bytes() + model.lm_head.weight_bytes()) / 1e9,
                             model.kv_read_bytes([a.ctx] * a.batch) / 1e9)},
           "e2e": {"value": tok / t_e2e, "unit": "tok/s", "h2d_bytes_per_step": eng.h2d_bytes(hi),
                   "d2h_bytes_per_step": eng.h_out.numel() * eng.h_out.element_size(),
                   "ms_per_step": t_e2e / a.steps * 1e3},
           "gpu_launches": launches * a.steps, "gpu_launches_per_step": per_op, "clocks": clk}
    if parity is not None:
        out["parity_check"] = parity
    if a.layers is not None:
        out["config"]["INVALID"] = "truncated model (--layers)"
    if world > 1 and not a.no_extra_models and a.model == "llama-3-8b" and a.quant == "w4a16" and a.layers is None:
        # north_star asks for Llama-3-70B at TP 2/4/8: W4A16 g128 and W8A8 FP8 + fp8 KV, same batch / ctx, at this TP degree
        extra = {}
        del eng
        for label, quant, kv in (("llama-3-70b w4a16 g128, kv auto", "w4a16", "auto"),
                                 ("llama-3-70b w8a8 fp8 (cutlass_scaled_mm), dynamic per-tensor activation scale, kv fp8", "fp8", "fp8"),
                                 ("llama-3-70b w8a8 fp8 (cutlass_scaled_mm), static activation scales fused into norm/silu, kv fp8",
                                  "fp8_static", "fp8")):
            try:
                extra[label] = time_extra_model(a, decoder, decoder.LLAMA3_70B, quant, kv, dev, world, rank, model.custom_ar,
                                                max(5, a.steps // 2))
            except Exception as e:   # keep the headline line even if an extra model cannot run on this box
                extra[label] = {"error": repr(e)[:200]}
            torch.cuda.empty_cache()
        out["extra_models_at_this_tp"] = extra
    if rank == 0 and a.quant in ("w4a16", "w4a16_24"):
        out["roofline"] = gemm_roofline(model, a.batch, peaks)
        step_alg = (model.linear_weight_bytes() + model.lm_head.weight_bytes() + model.kv_read_bytes([a.ctx] * a.batch))
        out["step_hbm_floor_ms"] = step_alg / (peaks.get("hbm_gbs", 6650.0) * 1e9) * 1e3
    if not a.no_sweep and world == 1:
        # the metric is quoted over "bs 1-64" (benchmark_marlin.py:20-21 sweeps the same batches)
        sw = {str(a.batch): tok / t_dev}
        for b in (1, 2, 4, 8, 16, 32):
            if b == a.batch:
                continue
            e = decoder.DecodeEngine(model, b, a.ctx)
            e.load_inputs(e.host_inputs(a.ctx, seed=3))
            e.prepare(a.ctx)
            for _ in range(3):
                e.run_device()
            torch.cuda.synchronize()
            e0.record()
            for _ in range(a.steps):
                e.run_device()
            e1.record()
            torch.cuda.synchronize()
            sw[str(b)] = b * a.steps / (e0.elapsed_time(e1) * 1e-3)
            del e
        out["tok_s_by_batch"] = dict(sorted(sw.items(), key=lambda kv: int(kv[0])))
        if "roofline" in out:
            out["roofline"]["by_batch"] = {str(a.batch): {"frac": out["roofline"]["frac"], "avg_launch_us": out["roofline"]["avg_launch_us"]}}
            for b in (1, 16):
                if b != a.batch:
                    out["roofline"]["by_batch"][str(b)] = gemm_roofline(model, b, peaks, brief=True)
    if not a.no_sweep and world == 1 and a.layers is None:
        # BASELINE.json configs[2] names "decode + 2k prefill": one 2048-token prompt through the same model (M = 2048 GEMMs,
        # rotary, reshape_and_cache, prefill attention over the paged cache, last-token logits), device-timed, median of 5
        try:
            T = 2048
            toks = torch.randint(0, shape.vocab, (T,), device=dev)
            bt = torch.randperm(model.num_blocks)[:T // model.block_size].to(torch.int32).view(1, -1)
            pos, pmeta = model.prefill_meta([T], [0], bt)
            model.forward(toks, pos, pmeta)
            torch.cuda.synchronize()
            ts = []
            for _ in range(5):
                e0.record()
                model.forward(toks, pos, pmeta)
                e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            ts.sort()
            out["prefill_2k"] = {"tokens": T, "ms": ts[2], "tok_s": T / ts[2] * 1e3,
                                 "what": "1 x 2048-token prompt, eager launches, kv auto; not part of `value`"}
        except Exception as e:   # never lose the headline line to the extra
            out["prefill_2k"] = {"error": repr(e)[:200]}
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        from oracle import cpu_decode
        r = cpu_decode.time_decode(shape, a.batch, a.ctx, layer_reps=3, threads=os.cpu_count())
        out["cpu_baseline"] = {"value": a.batch / r["step_s"], "unit": "tok/s", "cores": r["cores"],
                               "host_cpu_count": os.cpu_count(), "kind": r["kind"], "sample": r["sample"]}
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        # The decode graph holds captured NCCL all-reduces: tearing the communicator down while the graph execs are
        # alive (destroy_process_group, or interpreter teardown in arbitrary order) blocks forever on B200 / NCCL 2.28.
        # Ranks meet once more (rank 0 may still have been timing the GEMMs alone), then leave without the destructors.
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
