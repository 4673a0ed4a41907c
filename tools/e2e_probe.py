import sys, torch
sys.path.insert(0, ".")
from nm_vllm_b200 import decoder
from oracle import decode_ref
dev = "cuda:0"
def run(mode):
    model = decoder.SyntheticLlama(decoder.TINY, device=dev, num_blocks=64, max_pos=256)
    eng = decoder.DecodeEngine(model, batch=4, max_ctx=96)
    hi = eng.host_inputs(80, seed=0)
    kv0 = [(k.cpu().clone(), v.cpu().clone()) for k, v in model.kv]
    if mode == "step_first":
        eng.step(hi); eng.step(hi)
    else:
        eng.load_inputs(hi)
    seen = {}
    orig = decoder.ops.greedy_sample
    def cap(l):
        seen["l"] = l.float().cpu(); return orig(l)
    decoder.ops.greedy_sample = cap
    eng._device_step(80); torch.cuda.synchronize()
    decoder.ops.greedy_sample = orig
    ref = decode_ref.decode_step_logits(model, hi, kv0)
    print(mode, "err", float((seen["l"] - ref).abs().mean() / ref.abs().mean()), "slots", hi["slot_mapping"].tolist())
run("eager_first"); run("step_first"); run("eager_first")
