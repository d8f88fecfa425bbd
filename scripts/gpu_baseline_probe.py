import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from torch.profiler import profile, ProfilerActivity
dev = torch.device("cuda", 0)
print(bench.gpu_baseline(dev, steps=8, prefix_frames=1290))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import zonos_oracle as zo
cfg = zo.OracleConfig(eos_token_id=1025)
sd = {k: v.to(dev) for k, v in zo.synth_state_dict(cfg, 0).items()}
prefix = zo.synth_prefix_conditioning(cfg, 1, 24, 1).to(dev)
delayed = zo.apply_delay_pattern(torch.randint(0, 1024, (1, 9, 1400)), 1025).to(dev)
with torch.inference_mode(), torch.device(dev):
    o = zo.OracleZonos(cfg, sd)
    params = o.setup_cache(2, 24 + 1400 + 18)
    o.prefill(prefix, delayed[..., :1291], params, 2.0)
    params.seqlen_offset += 24 + 1291
    params.lengths_per_sample[:] += 24 + 1291
    for s in range(3):
        o.decode_one_token(delayed[..., 1291 + s: 1292 + s], params, 2.0)
        params.seqlen_offset += 1; params.lengths_per_sample[:] += 1
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        o.decode_one_token(delayed[..., 1295:1296], params, 2.0)
        torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=14))
