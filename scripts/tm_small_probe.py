"""Probe: small-config generate through the tensor-memory-ring instantiation, one variant per process (a hang is a timeout)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import zonos_oracle as zo
import zonos_vibes_b200 as z
spp, ntok, opts = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3:]
cfg = zo.SMALL_CONFIG
dev = torch.device("cuda", 0)
m = z.Zonos.from_state_dict(z.ZonosConfig.from_dict(cfg.to_reference_dict()), zo.synth_state_dict(cfg, 0), dev)
ctx = m._native()
for kv in opts:
    k, v = kv.split("="); ctx.set_option(k, int(v))
prefix = zo.synth_prefix_conditioning(cfg, 1, 6, 1).to(dev)
out = m.generate(prefix, max_new_tokens=ntok, sampling_params=dict(temperature=0), progress_bar=False, steps_per_poll=spp)
torch.cuda.synchronize()
print("ok", spp, ntok, opts, tuple(out.shape), flush=True)
