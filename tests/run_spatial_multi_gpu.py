"""Multi-GPU check of the spatial mode (run under torchrun on N GPUs of one box):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29533 tests/run_spatial_multi_gpu.py [--h 2160 --w 3840 --n 200000 --iters 20]
Every rank fits the same frame on its band of tile rows; the gradient rows are summed across ranks
either by one NCCL all-reduce per iteration (--exchange nccl) or inside the update kernel through
peer-mapped memory (--exchange peer); rank 0 also runs the whole-frame fit alone and compares losses / parameters,
then prints one JSON line with the timing of both (CUDA events, max over ranks)."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--h", type=int, default=2160)
    ap.add_argument("--w", type=int, default=3840)
    ap.add_argument("--n", type=int, default=200_000)
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--exchange", default="nccl", choices=["nccl", "peer", "nvls", "sharded"])
    ap.add_argument("--phase-times", type=int, default=0,
                    help="after the timed run: this many extra iterations with CUDA events around every phase call")
    ap.add_argument("--kernel-table", type=int, default=0,
                    help="after everything else: this many iterations under torch.profiler, rank 0 prints the "
                         "per-kernel device times (diagnosis only; never a reported figure)")
    ap.add_argument("--torch-barrier", action="store_true",
                    help="peer exchanges: torch's symmetric-memory barrier between two host calls instead of the library's own")
    ap.add_argument("--no-check", action="store_true", help="timing of a deliberately broken kernel variant")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from gsvc_b200 import synth
    from gsvc_b200.fit import FusedFrameFit, SpatialFrameFit

    init = synth.gaussians(a.n, seed=1)
    gt = (synth.target_frame(a.h, a.w)[0] * 255).round().to(torch.uint8).permute(1, 2, 0).contiguous().to(dev)
    fit = SpatialFrameFit(a.h, a.w, a.n, dev, init=init, exchange=a.exchange, own_barrier=not a.torch_barrier)
    fit.train_iters(gt, a.warmup)
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fit.train_iters(gt, a.iters)
    e1.record()
    dist.barrier(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / a.iters], dtype=torch.float64, device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    loss = fit.loss()
    # all ranks must hold bit-identical parameters
    chk = torch.stack([fit._xyz.double().sum(), fit._cholesky.double().sum(), fit._features_dc.double().sum()])
    lo, hi = chk.clone(), chk.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    identical = bool(torch.equal(lo, hi))
    out = None
    if rank == 0:
        ref = FusedFrameFit(a.h, a.w, a.n, dev, init=init)
        ref.train_iters(gt, a.warmup)
        torch.cuda.synchronize()
        e0.record()
        ref.train_iters(gt, a.iters)
        e1.record()
        torch.cuda.synchronize()
        ref_ms = e0.elapsed_time(e1) / a.iters
        ref_loss = ref.loss()
        d = (ref._features_dc - fit._features_dc).abs()
        out = {"world": world, "exchange": a.exchange, "H": a.h, "W": a.w, "gaussians": a.n, "iters": a.iters, "bands": fit.bands,
               "spatial_ms_per_iter": float(ms), "single_gpu_ms_per_iter": ref_ms, "speedup": ref_ms / float(ms),
               "loss_spatial": loss, "loss_single": ref_loss, "params_identical_across_ranks": identical,
               "median_abs_param_diff_vs_single": float(d.median()), "exchange_bytes_per_rank": a.n * 48}
        assert identical or a.no_check, "ranks diverged"
        assert abs(loss - ref_loss) <= 2e-3 * ref_loss or a.no_check, (loss, ref_loss)
    phase_ms = None
    dist.barrier()
    if a.phase_times:
        # where an iteration goes: events around every gs2d_fit_step_band call; what lies between two
        # calls is the barrier / collective.  Not part of the timed figure above.
        marks, call = [], fit._band_call

        def timed(gt_, fmt, band, phases, *rest, **kw):
            e_in, e_out = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e_in.record()
            call(gt_, fmt, band, phases, *rest, **kw)
            e_out.record()
            marks.append((phases, e_in, e_out))

        fit._band_call = timed
        fit.train_iters(gt, a.phase_times)
        torch.cuda.synchronize()
        acc = {}
        for k, (ph, e_in, e_out) in enumerate(marks):
            acc[f"phase{ph}"] = acc.get(f"phase{ph}", 0.0) + e_in.elapsed_time(e_out)
            if k + 1 < len(marks):
                acc[f"after{ph}"] = acc.get(f"after{ph}", 0.0) + e_out.elapsed_time(marks[k + 1][1])
        t = torch.tensor([acc[k] / a.phase_times for k in sorted(acc)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        phase_ms = dict(zip(sorted(acc), [round(float(x), 4) for x in t]))
        fit._band_call = call
    if rank == 0:
        out["phase_ms_max_over_ranks"] = phase_ms
        print(json.dumps(out), flush=True)
    if a.kernel_table:
        from torch.profiler import ProfilerActivity, profile
        dist.barrier()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            fit.train_iters(gt, a.kernel_table)
            torch.cuda.synchronize()
        rows = [(e.key, e.count, e.device_time_total / max(e.count, 1)) for e in prof.key_averages()]
        text = "".join(f"KERNEL rank {rank} {us:9.2f} us x{cnt:5d}  {key[:100]}\n"
                       for key, cnt, us in sorted(rows, key=lambda r: -r[1] * r[2]) if us > 0)
        for r in range(world):        # one rank at a time, so that the lines do not interleave
            if r == rank:
                print(text, end="", flush=True)
            dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
