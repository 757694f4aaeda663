#!/usr/bin/env python
"""Benchmark of the 2D-3D matching hot path (BASELINE.json: "2D-3D matched frames/sec").

Workload (BASELINE.json configs[1]): spot_demo-shaped synthetic sequence -- 512x512 query frames against one
8k-point 3D model, random-init weights of the reference architecture (calibrated + planted so that matches
exist, onepose_st_b200/synthetic.py), frames sharded round-robin over the GPUs (weak scaling: every rank
processes ``--frames-per-step`` frames per step), matches gathered once at the end over NCCL.

One JSON line on stdout (rank 0).  ``value`` = frames/s with frames resident in HBM; ``e2e`` = frames/s through
``OnePosePlus_model.forward(data)`` with the frame copied from pinned host memory and the matches copied back
every forward.  ``roofline`` = the sweep kernel of CoarseMatching (sim_fast_kernel, the kernel north_star grades), timed live by
its own CUDA-event pair on the launching stream inside the timed region (``roofline.op_level`` = the whole
opb200_coarse_match call); ``stages`` / ``roofline_other`` give the call-level figures of every other op.
``--impl reference`` times the reference algorithm's CPU port (oracle/, torch CPU fp32, all host threads) on
the same config -- the only place besides ``cpu_baseline`` where bench.py executes oracle code.
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
if REPO not in sys.path:
    sys.path.insert(0, REPO)

METRIC = "2D-3D matched frames/sec"
UNIT = "frames/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--points", type=int, default=8000, help="3D points of the object model (L)")
    ap.add_argument("--size", type=int, default=512, help="query crop side (pixels)")
    ap.add_argument("--frames-per-step", type=int, default=16, help="frames each rank processes per step")
    ap.add_argument("--batch", type=int, default=8, help="frames per forward call (dict contract N)")
    ap.add_argument("--pool", type=int, default=500, help="frames of the synthetic sequence")
    ap.add_argument("--plant", type=int, default=1000, help="planted 2D-3D correspondences")
    ap.add_argument("--precision", type=int, default=3, choices=[1, 3], help="3 = bf16x3 split (fp32-grade), 1 = bf16")
    ap.add_argument("--cpu-sample-frames", type=int, default=4, help="frames timed for cpu_baseline (0 = skip)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--cuda-graph", type=int, default=1, help="replay the forward's head (backbone .. coarse matching) as a CUDA graph")
    ap.add_argument("--streams", type=int, default=3, help="host threads x CUDA streams per GPU that pull frame batches")
    return ap.parse_args()


def load_peaks():
    p = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU every 20 ms through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples, self.reasons, self.sm_max = [], set(), None
        self._stop_ev = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40,
                 "hw_power_brake_slowdown": 0x80, "sw_power_cap": 0x4, "sync_boost": 0x10}
        while not self._stop_ev.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for n, bit in names.items():
                    if r & bit:
                        self.reasons.add(n)
            except Exception:
                pass
            self._stop_ev.wait(0.02)

    def stop(self):
        self._stop_ev.set()
        self.join(timeout=2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.sm_max, "reasons": sorted(self.reasons), "samples": 0}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.sm_max, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def config_dict(args, n_gpus):
    return {"workload": f"spot_demo-shaped synthetic sequence: {args.size}x{args.size} query frames vs "
                        f"{args.points}-point 3D model (BASELINE.json configs[1])",
            "points": args.points, "image": [args.size, args.size], "coarse_grid": [args.size // 8, args.size // 8],
            "frames_per_step_per_gpu": args.frames_per_step, "frames_per_forward": args.batch,
            "planted_matches": args.plant, "precision": "bf16x3-split (fp32-grade)" if args.precision == 3 else "bf16",
            "sharding": f"frames round-robin over {n_gpus} GPU(s); one NCCL gather of matches at the end",
            "host_pipeline": f"{args.streams} host threads x CUDA streams per GPU pull frame batches from one queue "
                             "(onepose_st_b200.pipeline.StreamPipeline)",
            "cuda_graph": bool(args.cuda_graph),
            "l2": "256 MiB buffer written between steps (inside the timed region)"}


# ------------------------------------------------------------------------------------------------ CPU reference
def cpu_reference_setup(args):
    """Weights, 3D model and frames for the CPU port -- built on CPU with the same seeds/recipe as the GPU arm."""
    import torch
    from onepose_st_b200 import synthetic
    from onepose_st_b200.onepose_plus import OnePosePlus_model
    torch.manual_seed(0)
    cfg = synthetic.onepose_plus_cfg()
    model = OnePosePlus_model(cfg).eval()
    sd = synthetic.calibrate_onepose_state_dict(model.state_dict())
    model.load_state_dict(sd, strict=True)
    seq = synthetic.make_onepose_sequence(model, args.points, 8, args.size, args.size, args.plant, device=torch.device("cpu"))
    return cfg, sd, seq


def cpu_reference_frame(R, cfg, sd, seq, f):
    import torch
    data = {"keypoints3d": seq["keypoints3d"], "descriptors3d_db": seq["descriptors3d_db"],
            "descriptors3d_coarse_db": seq["descriptors3d_coarse_db"], "query_image": seq["frames"][f % 8][None]}
    with torch.no_grad():
        R.onepose_plus_forward(sd, cfg, data)
    return int(data["mconf"].shape[0])


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    from oracle import ref_ops as R  # the reference algorithm's CPU port (bench.py may execute it here only)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg, sd, seq = cpu_reference_setup(args)
    for w in range(args.warmup):
        cpu_reference_frame(R, cfg, sd, seq, w)
    t0 = time.perf_counter()
    m = 0
    for s in range(args.steps):
        m += cpu_reference_frame(R, cfg, sd, seq, s)
    dt = time.perf_counter() - t0
    fps = args.steps / dt
    sample = f"1 frame per step ({args.steps} frames timed), N=1 forward, torch CPU fp32, {torch.get_num_threads()} threads"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config_dict(args, args.gpus),
        "cpu_baseline": {"value": fps, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port", "sample": sample},
        "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "matches_per_frame": m / max(1, args.steps), "gpu_launches": 0}))


# ------------------------------------------------------------------------------------------------ GPU arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    import threading
    from onepose_st_b200 import _lib, ops, pipeline, sharding, synthetic
    from onepose_st_b200.onepose_plus import OnePosePlus_model

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: onepose_st_b200 has no CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # NCCL's banner / warnings off stdout: rank 0 prints ONE JSON line
        dist.init_process_group("nccl", device_id=dev)
    peaks = load_peaks()

    torch.manual_seed(0)
    cfg = synthetic.onepose_plus_cfg()
    model = OnePosePlus_model(cfg).eval()
    model.load_state_dict(synthetic.calibrate_onepose_state_dict(model.state_dict()), strict=True)
    model = model.to(dev)
    model.coarse_matching.precision = args.precision
    model.loftr_coarse.precision = args.precision
    model.loftr_fine.precision = args.precision
    seq = synthetic.make_onepose_sequence(model, args.points, args.pool, args.size, args.size, args.plant, device=dev)
    frames_dev = seq["frames"]
    frames_host = frames_dev.cpu().pin_memory()
    B, F = args.batch, args.frames_per_step
    assert F % B == 0, "--frames-per-step must be a multiple of --batch"
    kp = seq["keypoints3d"].expand(B, -1, -1)  # one 3D model for every frame of the batch: broadcast views
    d_f = seq["descriptors3d_db"].expand(B, -1, -1)
    d_c = seq["descriptors3d_coarse_db"].expand(B, -1, -1)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    tl = threading.local()  # per host thread: device staging buffer of the end-to-end arm
    ragged = sharding.RaggedRows(6, dev, capacity_rows=1 << 18)
    n_units_total = world * F * (args.steps + args.warmup)

    def unit_frames(step, b0):
        # global unit id -> frame of the pool; rank r owns units r, r+G, ...
        return [((step * F + b0 + i) * world + rank) % args.pool for i in range(B)]

    # device-side frame indices of every forward (value arm): addressing only, prepared before the timed region
    idx_dev = {(s_, b0): torch.tensor(unit_frames(s_, b0), device=dev)
               for s_ in range(args.steps + args.warmup) for b0 in range(0, F, B)}

    def forward(images):
        data = {"keypoints3d": kp, "descriptors3d_db": d_f, "descriptors3d_coarse_db": d_c, "query_image": images}
        model(data)
        return data

    def one_forward(job):
        """One forward of B frames on the calling thread's current stream; returns (rows [M,6], per-frame counts, D2H bytes)."""
        s, b0, e2e = job
        if b0 == 0:
            flush.fill_(s & 0xFF)  # L2 flush between steps
        idx = unit_frames(s, b0)
        nbytes = 0
        if e2e:
            stage = getattr(tl, "stage", None)
            if stage is None:
                stage = tl.stage = torch.empty(B, 1, args.size, args.size, device=dev)
            for i, f in enumerate(idx):
                stage[i].copy_(frames_host[f], non_blocking=True)
            data = forward(stage)
            rows = sharding.rows_onepose(data)
            host_rows = rows.cpu()  # D2H of the step's result (synchronises this stream)
            nbytes = host_rows.numel() * 4
        else:
            data = forward(frames_dev.index_select(0, idx_dev[(s, b0)]))
            rows = sharding.rows_onepose(data)
        counts = torch.bincount(data["b_ids"], minlength=B).tolist() if B > 1 else [rows.shape[0]]
        return rows, counts, nbytes

    def run_steps(first, count, e2e, pipe):
        jobs = [(s, b0, e2e) for s in range(first, first + count) for b0 in range(0, F, B)]
        d2h = 0
        for rows, counts, nbytes in pipe.map(one_forward, jobs):
            rows.record_stream(torch.cuda.current_stream())
            off = 0
            for c in counts:
                ragged.append(rows[off: off + c])
                off += c
            d2h += nbytes
        return d2h

    pipes = {}

    def timed(e2e, n_streams, stage_events=False):
        pipe = pipes.get(n_streams)
        if pipe is None:
            pipe = pipes[n_streams] = pipeline.StreamPipeline(n_streams, dev)
        # CUDA graph of the forward's head (one launch instead of ~85); the per-op event pass runs eagerly
        model.cuda_graph = bool(args.cuda_graph) and not stage_events
        ragged.reset()
        one_forward((0, 0, e2e))  # builds the per-model caches (packed weights, encoded 3D model) on one stream
        torch.cuda.synchronize()
        pipe.warm(lambda: one_forward((0, B if F > B else 0, e2e)))  # per-stream workspaces / CUDA graphs
        run_steps(0, args.warmup, e2e, pipe)
        # the gather is warmed up too (first NCCL all_gather sets up channels; first sort/scatter compiles nothing but allocates)
        # at the size of the timed region (the warm-up rows repeated), so that the timed gather finds its buffers in the
        # caching allocator instead of calling cudaMalloc
        rep_ = max(1, -(-args.steps // max(1, args.warmup))) + 1  # a little larger: bigger cached blocks serve smaller requests
        sharding.gather_matches(ragged.rows().repeat(rep_, 1), list(ragged.counts) * rep_, world * F * args.warmup * rep_)
        # ... and once more at (nearly) exactly the timed shape: the first F * steps units of the repeated rows, so that
        # NCCL meets the timed message size and the allocator the timed block sizes before the clock starts
        cl_ = (list(ragged.counts) * rep_)[: F * args.steps]
        sharding.gather_matches(ragged.rows().repeat(rep_, 1)[: sum(cl_)], cl_, world * len(cl_))
        ragged.reset()
        ops.stage_events = {} if stage_events else None
        sampler = ClockSampler(local_rank)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        sampler.start()
        l0 = _lib.launch_count()
        mallocs0 = torch.cuda.memory_stats(dev).get("num_device_alloc", 0)
        ev0, ev1, evg = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        if stage_events:
            _lib.coarse_kernel_ms()  # forget the sweep-kernel brackets of the warm-up launches
        ev0.record()
        d2h = run_steps(args.warmup, args.steps, e2e, pipe)
        n_units = world * F * args.steps
        evg.record()
        rows_all, counts_all = sharding.gather_matches(ragged.rows(), ragged.counts, n_units)
        ev1.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        launches = _lib.launch_count() - l0
        clocks = sampler.stop()
        ms = torch.tensor([ev0.elapsed_time(ev1)], device=dev)
        per_rank = [float(ms.item())]
        if world > 1:
            allms = torch.empty(world, device=dev)
            dist.all_gather_into_tensor(allms, ms)
            per_rank = [round(v, 2) for v in allms.tolist()]
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        events, ops.stage_events = ops.stage_events, None
        diag = {"ms_per_rank": per_rank, "gather_ms_rank0": round(evg.elapsed_time(ev1), 3),
                "cuda_mallocs_in_timed_region": torch.cuda.memory_stats(dev).get("num_device_alloc", 0) - mallocs0}
        return {"ms": float(ms.item()), "launches": launches, "clocks": clocks, "events": events, "d2h": d2h,
                "matches": int(rows_all.shape[0]), "units": n_units, "diag": diag}

    res = timed(e2e=False, n_streams=args.streams)
    res_e2e = None if args.no_e2e else timed(e2e=True, n_streams=args.streams)
    # per-op device time: the same K steps once more on ONE stream, so that the CUDA-event bracket around each C-ABI
    # call (recorded on the launching stream, inside this timed region) holds that call's kernels only -- with two
    # streams in flight a bracket would also span the other stream's kernels.
    _lib.coarse_kernel_timing(True)  # + an event bracket around the sweep kernel itself, inside opb200_coarse_match
    res_ops = timed(e2e=False, n_streams=1, stage_events=True)
    _lib.coarse_kernel_timing(False)
    sim_ms, sim_n = _lib.coarse_kernel_ms()

    frames_rank = F * args.steps
    L, S, C = args.points, (args.size // 8) ** 2, 256
    stages = {}
    for name, evs in (res_ops["events"] or {}).items():
        tot = sum(a.elapsed_time(b) for a, b in evs)
        stages[name] = {"calls": len(evs), "ms_per_frame": tot / frames_rank}
    m_per_frame = res["matches"] / max(1, res["units"])
    rooflines = {}
    traffic = None  # DRAM bytes per launch of sim_fast_kernel, from the committed ncu --set full capture (profiles/r01_coarse_match_ncu_full_summary.txt)
    tpath = os.path.join(REPO, "profiles", "r01_traffic.json")
    if os.path.exists(tpath) and args.batch == 8 and args.points == 8000 and args.size == 512:
        with open(tpath) as f:
            traffic = json.load(f)["sim_fast_kernel"]["dram_bytes_per_launch"]
    if "coarse_match" in stages:
        t_op = stages["coarse_match"]["ms_per_frame"] * 1e-3
        flops = 2.0 * L * S * C
        peak = peaks["bf16_tflops_sustained"]
        # the dominant kernel of CoarseMatching: sim_fast_kernel, timed by its own event pair on the launching stream
        # (opb200_coarse_kernel_timing) inside the timed region of the single-stream pass
        n_timed = stages["coarse_match"]["calls"]
        t_k = (sim_ms / sim_n) * 1e-3 if sim_n else None  # all recorded launches have the same shape
        executed = 1  # the sweep runs ONE fp16 x fp16 product; captured elements are re-evaluated from the split planes
        op = {"achieved": flops / t_op / 1e12, "frac": flops / t_op / 1e12 / peak, "avg_ms_per_call": t_op * 1e3 * B,
              "note": "the whole opb200_coarse_match call: fp16 operand pack, sweep, partial merge, exact re-evaluation of the "
                      "captured elements, mutual-NN resolve, ordered compaction"}
        if t_k is not None:
            ach = flops * B / t_k / 1e12
            rooflines["coarse_match"] = {
                "bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak, "traffic": traffic,
                "launch": {"frames_per_launch": B, "algorithmic_flop_per_launch": flops * B, "avg_ms_per_launch": t_k * 1e3,
                           "launches_timed": sim_n, "calls_in_timed_steps": n_timed},
                "kernel": "sim_fast_kernel<true>: single-pass dual-softmax sweep of CoarseMatching (tcgen05 fp16 similarity tiles, "
                          "row/column softmax sums and candidate capture in the epilogue)",
                "algorithmic": f"2*L*S*C = {flops / 1e9:.1f} GFLOP per frame (one similarity GEMM)",
                "executed_flop_multiple": executed, "peak_source": peaks["source"] + " bf16 sustained", "op_level": op}
        else:  # the two-pass path ran (masks / conf matrix requested): only the call-level figure exists
            rooflines["coarse_match"] = {"bound": "tensor", "achieved": op["achieved"], "peak": peak, "unit": "TFLOP/s",
                                         "frac": op["frac"], "traffic": None, "kernel": "opb200_coarse_match (two-pass path)",
                                         "algorithmic": f"2*L*S*C = {flops / 1e9:.1f} GFLOP per frame", "op_level": op}
    if "transformer256" in stages:
        t = stages["transformer256"]["ms_per_frame"] * 1e-3
        flops = 120.0 * (L + S) * C * C
        peak = peaks["bf16_tflops_sustained"]
        rooflines["transformer256"] = {"bound": "tensor", "achieved": flops / t / 1e12, "peak": peak, "unit": "TFLOP/s",
                                       "frac": flops / t / 1e12 / peak, "traffic": None,
                                       "algorithmic": f"120*(L+S)*C^2 = {flops / 1e9:.1f} GFLOP per frame",
                                       "executed_flop_multiple": (3 if args.precision == 3 else 1)}
    if "backbone_conv" in stages:
        t = stages["backbone_conv"]["ms_per_frame"] * 1e-3
        h2, h4, h8 = args.size // 2, args.size // 4, args.size // 8
        flops = 2.0 * (h2 * h2 * (4 * 9 * 128 * 128 + 9 * 196 * 196 + 9 * 196 * 128 + 128 * 196)
                       + h4 * h4 * (9 * 128 * 196 + 128 * 196 + 3 * 9 * 196 * 196 + 9 * 256 * 256 + 9 * 256 * 196 + 196 * 256)
                       + h8 * h8 * (9 * 196 * 256 + 196 * 256 + 3 * 9 * 256 * 256 + 256 * 256))
        peak = peaks["bf16_tflops_sustained"]
        rooflines["backbone_conv"] = {"bound": "tensor", "achieved": flops / t / 1e12, "peak": peak, "unit": "TFLOP/s",
                                      "frac": flops / t / 1e12 / peak, "traffic": None,
                                      "kernel": "conv_kernel: fp16-operand tcgen05 implicit GEMM, 23 convolutions of ResNetFPN_8_2",
                                      "algorithmic": f"sum 2*Ho*Wo*k*k*Cin*Cout = {flops / 1e9:.1f} GFLOP per frame",
                                      "executed_flop_multiple": 1}
    if "fine_match" in stages and m_per_frame > 0:
        t = stages["fine_match"]["ms_per_frame"] * 1e-3
        by = m_per_frame * (4 * 128 * (25 + 1) + 20)
        rooflines["fine_match"] = {"bound": "hbm", "achieved": by / t / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                                   "frac": by / t / 1e9 / peaks["hbm_gbs"], "traffic": None,
                                   "algorithmic": "4*Cf*(WW+1)+20 bytes per match"}

    out = None
    if rank == 0:
        total_frames = world * F * args.steps
        value = total_frames / (res["ms"] * 1e-3)
        out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": res["ms"] / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
               "dtype": ("bf16x3 split products (matcher), " if args.precision == 3 else "bf16 (matcher), ")
               + ("fp16 operands (backbone convolutions, TF32-grade like the reference on a GPU), "
                  if torch.backends.cudnn.allow_tf32 else "fp32 cuDNN backbone, ") + "f32 accumulate/IO",
               "data": "synthetic", "config": config_dict(args, world),
               "matches_per_frame": m_per_frame, "gpu_launches": res["launches"], "clocks": res["clocks"],
               "streams": args.streams, "diag": res["diag"],
               "single_stream": {"value": total_frames / (res_ops["ms"] * 1e-3), "unit": UNIT,
                                 "note": "same steps on one CUDA stream; stages_ms_per_frame and the rooflines are timed in this pass"},
               "stages_ms_per_frame": {k: round(v["ms_per_frame"], 4) for k, v in stages.items()},
               "roofline": rooflines.get("coarse_match"),
               "roofline_other": {k: v for k, v in rooflines.items() if k != "coarse_match"}}
        if res_e2e is not None:
            out["e2e"] = {"value": total_frames / (res_e2e["ms"] * 1e-3), "unit": UNIT,
                          "h2d_bytes_per_step": F * args.size * args.size * 4,
                          "d2h_bytes_per_step": res_e2e["d2h"] // max(1, args.steps),
                          "ms_per_step": res_e2e["ms"] / args.steps, "gpu_launches": res_e2e["launches"],
                          "clocks": res_e2e["clocks"], "diag": res_e2e["diag"],
                          "api": "OnePosePlus_model.forward(data): pinned-host frame -> H2D -> forward -> D2H of [M,6] match rows"}
    # CPU baseline: the reference algorithm's CPU port on this box's host cores (rank 0, N=1 only)
    if rank == 0 and world == 1 and args.cpu_sample_frames > 0:
        from oracle import ref_ops as R
        torch.set_num_threads(os.cpu_count() or 1)
        ccfg, csd, cseq = cpu_reference_setup(args)
        cpu_reference_frame(R, ccfg, csd, cseq, 0)
        t0 = time.perf_counter()
        for f in range(args.cpu_sample_frames):
            cpu_reference_frame(R, ccfg, csd, cseq, f + 1)
        dt = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": args.cpu_sample_frames / dt, "unit": UNIT, "cores": torch.get_num_threads(),
                               "kind": "port",
                               "sample": f"{args.cpu_sample_frames} frames of the same workload after 1 warm-up frame, "
                                         "N=1 forward each, torch CPU fp32"}
    elif rank == 0:
        out["cpu_baseline"] = None
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
