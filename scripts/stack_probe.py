"""Forward of a 16-layer fused encoder stack at config-3 width, per layer, with and without the op-chain kernel."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pipe_b200.models.fused import FusedEncoderLayer
from pipe_b200.partition import PipeSequential

dev = torch.device("cuda:0")
E, H, FF, S, NB, L = 4096, 16, 4096, 25, 3, int(os.environ.get("LAYERS", 16))
torch.manual_seed(0)
layers = [FusedEncoderLayer(E, H, FF, 0.2).to(dev) for _ in range(L)]
stack = PipeSequential(*layers)
x = torch.randn(S, NB, E, device=dev).bfloat16()

def run():
    with torch.no_grad():
        return stack(x)

def time_us(n=20, warm=5):
    for _ in range(warm): run()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): run()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / n

outs = {}
for chain in ("0", "1", "0", "1"):
    os.environ["PB2_CHAIN"] = chain
    torch.manual_seed(5)
    outs[chain] = run().float()
    t = time_us()
    print(f"PB2_CHAIN={chain}: {t:.1f} us per forward of {L} layers = {t / L:.1f} us per layer", flush=True)
print("rel diff chain vs separate:", ((outs["1"] - outs["0"]).norm() / outs["0"].norm()).item())
