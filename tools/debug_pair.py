"""GPU debug aid: run one GEMM shape (argv: M N K) and print the result or the CUDA error."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
os.environ["CUDA_LAUNCH_BLOCKING"] = "1"
import torch
from fangyan_tts_b200.engine import Engine
m, n, k = (int(x) for x in sys.argv[1:4])
eng = Engine(0, "bf16x3")
torch.manual_seed(0)
a = torch.randn(m, k, device="cuda").bfloat16()
b = torch.randn(n, k, device="cuda").bfloat16()
print("launch", (m, n, k), flush=True)
try:
    got = eng.gemm_bf16(a, b)
    torch.cuda.synchronize()
    ref = a.float() @ b.float().T
    print((m, n, k), "ok, rel err", float((got - ref).abs().max() / ref.abs().max()), flush=True)
except Exception as e:
    print((m, n, k), "FAILED:", repr(e)[:400], flush=True)
