"""The peer-memory filter step with a world of ONE rank (all four p2p kernels, exchanges with no peer) beside the
single-GPU step: what the p2p kernel variants cost by themselves, without any cross-GPU wait.
  python tools/diag_c3_world1.py [particles] [--p2p-only]"""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
import torch.distributed as dist

import bench_configs as bc


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    P = int(args[0]) if args else 1 << 20
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29571")
    torch.cuda.set_device(0)
    dist.init_process_group("gloo", rank=0, world_size=1)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    ctx = bc.Ctx(1, 0, 0, stream)
    H = bc.hmm_golden()
    T = len(H["data"])
    for sharded in ((True,) if "--p2p-only" in sys.argv else (False, True, False, True)):
        ms, tvd, _ = bc._c3_filter(ctx, H, P, sharded, 5, False)
        print("P = %d  %s: %.2f us/step (tvd %.4f)" % (P, "p2p kernels, world 1" if sharded else "single-GPU kernels  ", 1e3 * ms / T, tvd), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
