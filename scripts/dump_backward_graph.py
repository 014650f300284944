"""Print the autograd graph Pipe builds for a 2-stage, 2-micro-batch CPU pipeline — the structure the
reference's slides draw by hand (reference .pptx slides 1-3).

    python scripts/dump_backward_graph.py [never|always|except_last]

Every line is one autograd node, indented by its depth below the output's grad_fn; nodes reached a
second time are printed as a back-reference.  Fork / Join pairs are what force micro-batch i-1's
backward to run after micro-batch i's on the same stage; Checkpoint / Recompute are the two halves
of activation checkpointing; Copy / Wait are the stage-boundary transfers (no-ops on CPU streams,
but the same nodes order the real copies on GPUs)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch import nn

from pipe_b200 import Pipe


def dump(mode: str, out=sys.stdout) -> list:
    torch.manual_seed(0)
    model = nn.Sequential(nn.Linear(4, 4), nn.Tanh(), nn.Linear(4, 2))
    pipe = Pipe(model, chunks=2, checkpoint=mode, balance=[2, 1], devices=["cpu", "cpu"], strict_except_last=True)
    y = pipe(torch.rand(4, 4, requires_grad=True)).local_value()
    seen, names = {}, []

    def walk(fn, depth):
        if fn is None:
            return
        name = type(fn).__name__
        if fn in seen:
            print("  " * depth + f"^ {name} #{seen[fn]}", file=out)
            return
        seen[fn] = len(seen)
        names.append(name)
        print("  " * depth + f"{name} #{seen[fn]}", file=out)
        for nxt, _ in fn.next_functions:
            walk(nxt, depth + 1)

    walk(y.grad_fn, 0)
    pipe.close()
    return names


if __name__ == "__main__":
    dump(sys.argv[1] if len(sys.argv) > 1 else "except_last")
