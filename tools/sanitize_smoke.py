"""Small shapes of every kernel family in one process, each checked by its own bounds asserts -- a quick smoke
run (python tools/sanitize_smoke.py).  It was written as the target of a memory / race checker; this GPU pool refuses
such tools for good (profiles/r02_sanitizer_refused.log), so it stays as a plain smoke test.  Covered:

the three operand-major forms of the tcgen05 contraction, one single-pass and one recompute micro-batch through
PathStep (mbarrier / TMEM / cross-CTA protocols, lockstep counters), the K3 sort + segmented reduce, and the
world-1 form of the three peer exchanges (flag protocol on the own buffer; the 2-rank form runs in
tests/test_gpu_dp_peer.py)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "speech-integration_b200"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from ssi_b200 import ops, synth  # noqa: E402
from ssi_b200.step import PathStep  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    # the shared mainloop in its three operand-major forms: (K,K), (MN,MN), (K,MN)
    A = torch.randn(384, 320, device=dev).to(torch.bfloat16)
    B = torch.randn(520, 320, device=dev).to(torch.bfloat16)
    ops.debug_gemm(A, B)
    ops.debug_gemm(A.t().contiguous(), B.t().contiguous(), True, True)
    ops.debug_gemm(A, B.t().contiguous(), False, True)
    inp = synth.make_inputs(synth.CONFIGS["c2"], batch=2, seq=192, vocab=5002, hidden=256)
    args = [inp[k].to(dev) for k in ("tokens", "labels", "X", "dH0")]
    for mode in ("single_pass", "recompute"):
        step = PathStep(inp["W"].to(dev), mode=mode)
        step(*args)
        step.forward_only(args[2], args[1])
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:29617", rank=0, world_size=1)
    step = PathStep(inp["W"].to(dev), process_group=dist.group.WORLD, exchange="peer")
    step(*args)
    step(*args)
    torch.cuda.synchronize()
    step.close()
    dist.destroy_process_group()
    print("sanitize_smoke ok")


if __name__ == "__main__":
    main()
