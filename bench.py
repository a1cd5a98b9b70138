#!/usr/bin/env python
"""bench.py -- headline benchmark of the hot path (BASELINE.json):

  metric   GN-rays iterations/s on the 64-keyframe, 512x384 graph (63 chain + 65 loop
           closures, both directions -> E = 256 directed edges), edges sharded over N GPUs;
  matching px-matches/s of iter_proj + refine_matches on B frame pairs at 512x384
           (reported in the "matching" object of the same JSON line).

A "step" is one gauss_newton_rays call of `--gn-iters` (10, config/base.yaml:56) iterations
with delta_thresh = 0, i.e. a fixed iteration count.  `value` is timed with every input
resident in HBM; `e2e` is the same call through the reference-facing operator
(mast3r_slam_backends.gauss_newton_rays) with HOST (pinned) buffers: H2D of all inputs and
D2H of the poses inside the timed region.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
For N > 1 launch with torchrun (one rank per GPU).  See the task contract in DESIGN.md.
"""
import argparse
import contextlib
import io
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

H_IMG, W_IMG = 384, 512
N_PTS = H_IMG * W_IMG
GN_BYTES_PER_POINT = 45      # 1 + 8 + 12 + 12 + 4 + 4 + 4
MATCH_BYTES_PER_PX = 195     # 65 (iter_proj) + 130 (refine_matches fp16)


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index
        self.nvml_rows, self._stop, self._thread = [], threading.Event(), None

    # NVML is polled from a thread every 5 ms (the timed region of a default run is tens of
    # milliseconds: an `nvidia-smi -lms` child is too slow to start and to tick); nvidia-smi is
    # the fallback when the NVML binding cannot be used.
    def _nvml_loop(self, h, nv):
        bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown,
                "hw_thermal_slowdown": nv.nvmlClocksEventReasonHwThermalSlowdown,
                "sw_thermal_slowdown": nv.nvmlClocksEventReasonSwThermalSlowdown,
                "sw_power_cap": nv.nvmlClocksEventReasonSwPowerCap}
        mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        while not self._stop.is_set():
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                self.nvml_rows.append((float(sm), float(mx), [n for n, b in bits.items() if r & b]))
            except Exception:
                break
            time.sleep(0.005)

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            try:
                uuid = str(torch.cuda.get_device_properties(self.index).uuid)
                h = nv.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            except Exception:
                h = nv.nvmlDeviceGetHandleByIndex(self.index)
            nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
            self._thread = threading.Thread(target=self._nvml_loop, args=(h, nv), daemon=True)
            self._thread.start()
            return
        except Exception:
            self._thread = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-i", str(self.index), "-lms", "20"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
            t0 = time.time()
            while not self.rows and time.time() - t0 < 5.0:  # wait for the first tick
                time.sleep(0.02)
        except Exception:
            self.proc = None

    def mark(self):
        """Start of the timed region: earlier samples (warm-up) are dropped."""
        self.nvml_rows.clear()
        self.rows.clear()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self._thread is not None:
            self._stop.set()
            self._thread.join(timeout=1.0)
            sm = [r[0] for r in self.nvml_rows]
            reasons = sorted({n for r in self.nvml_rows for n in r[2]})
            return {"sm_mhz": float(np.median(sm)) if sm else None,
                    "sm_max_mhz": self.nvml_rows[0][1] if sm else None, "reasons": reasons,
                    "samples": len(sm), "source": "nvml, 5 ms period"}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.05)
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names)
                   if any(len(r) >= 6 and r[2 + k].lower() == "active" for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm), "source": "nvidia-smi -lms 20"}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def build_graph(device, num_kf_cfg="loop64"):
    from frodo_slam_b200 import synth
    K, und = synth.graph_edges(num_kf_cfg)
    g = synth.keyframe_graph(K, H_IMG, W_IMG, und, seed=1, device=device)
    return K, g


def timed_steps(fn, steps, warmup, barrier=None, on_start=None):
    """W warm-up calls, then K calls bracketed by barrier + synchronize; CUDA-event time (ms)."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if barrier:
        barrier()
    torch.cuda.synchronize()
    if on_start:
        on_start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    if barrier:
        barrier()
    return e0.elapsed_time(e1)


def gn_params():
    from frodo_slam_b200 import backends, synth
    gp = synth.GN_PARAMS
    return backends.make_params("rays", sigma_ray=gp["sigma_ray"], sigma_dist=gp["sigma_dist"],
                                C_thresh=gp["C_conf"], Q_thresh=gp["Q_conf"]), gp


def run_matching(device, B, steps, warmup, peak, be=None, hw=None):
    """px-matches/s of iter_proj + refine_matches over B pairs at 512x384 (or `hw`), inputs in
    HBM.  `be`: the backend module (default: this repo's drop-in; the reference arm passes the
    reference's own pybind module)."""
    if be is None:
        import mast3r_slam_backends as be
    from frodo_slam_b200 import synth
    mp = synth.MATCH_PARAMS
    H_IMG, W_IMG = hw if hw is not None else (globals()["H_IMG"], globals()["W_IMG"])
    N_PTS = H_IMG * W_IMG
    base = 8  # distinct synthetic pairs, tiled to B (inputs stay > L2: 38 MB/pair)
    d = synth.matching_pairs(min(base, B), H_IMG, W_IMG, seed=0, device=device)
    rep = (B + base - 1) // base
    tile = lambda x: x.repeat((rep,) + (1,) * (x.dim() - 1))[:B].contiguous()  # noqa: E731
    rays, pts, p0 = synth.prep_for_iter_proj(tile(d["X11"]), tile(d["X21"]))
    D11 = tile(d["D11"]).half().contiguous()
    D21 = tile(d["D21"]).half().reshape(B, N_PTS, -1).contiguous()
    del d
    out = {}

    def step():
        p, conv = be.iter_proj(rays, pts, p0, mp["max_iter"], mp["lambda_init"],
                               mp["convergence_thresh"])
        out["p1"] = p.long()
        out["r"] = be.refine_matches(D11, D21, out["p1"], mp["radius"], mp["dilation_max"])

    ms = timed_steps(step, steps, warmup)
    p1 = out["p1"]
    ms_ip = timed_steps(lambda: be.iter_proj(rays, pts, p0, mp["max_iter"], mp["lambda_init"],
                                             mp["convergence_thresh"]), steps, 1)
    ms_rm = timed_steps(lambda: be.refine_matches(D11, D21, p1, mp["radius"], mp["dilation_max"]),
                        steps, 1)
    px = B * N_PTS * steps
    val = px / (ms * 1e-3)
    return {
        "metric": "px-matches/s (iter_proj+refine_matches)", "value": val, "unit": "px/s",
        "pairs": B, "resolution": f"{W_IMG}x{H_IMG}", "ms_per_step": ms / steps,
        "iter_proj": {"px_per_s": px / (ms_ip * 1e-3),
                      "roofline_frac": 65 * px / (ms_ip * 1e-3) / 1e9 / peak},
        "refine_matches": {"px_per_s": px / (ms_rm * 1e-3),
                           "roofline_frac": 130 * px / (ms_rm * 1e-3) / 1e9 / peak,
                           # what actually bounds it: 5,880 packed-half
                           # instructions per pixel at one per 2 cycles per SM sub-partition
                           "bound": "packed-half pipe + shared-memory bandwidth",
                           "pipe_floor_px_per_s": 148 * 4 * 1.965e9 / 2 * 32 / 5880,
                           "frac_of_pipe_floor": px / (ms_rm * 1e-3) /
                                                 (148 * 4 * 1.965e9 / 2 * 32 / 5880)},
        "roofline": {"bound": "hbm", "achieved": MATCH_BYTES_PER_PX * val / 1e9, "peak": peak,
                     "unit": "GB/s", "frac": MATCH_BYTES_PER_PX * val / 1e9 / peak,
                     "traffic": None},
        "gpu_launches": 2 * steps,
    }


def run_match_pipeline(device, B, steps, warmup, peak):
    """px-matches/s through the drop-in ``match()`` (mast3r_slam/matching.py:8-10: prologue,
    iter_proj, occlusion test, fp16 cast, refine_matches, epilogue; one C-ABI call, four
    kernels), fp32 inputs resident in HBM, plus each HBM-bound glue kernel timed alone against
    its algorithmic bytes."""
    import ctypes
    from frodo_slam_b200 import _lib, matching as fm, synth
    from frodo_slam_b200.backends import _ptr, _stream
    base = min(8, B)
    d = synth.matching_pairs(base, H_IMG, W_IMG, seed=0, device=device)
    rep = (B + base - 1) // base
    tile = lambda x: x.float().repeat((rep,) + (1,) * (x.dim() - 1))[:B].contiguous()  # noqa: E731
    X11, X21, D11, D21 = tile(d["X11"]), tile(d["X21"]), tile(d["D11"]), tile(d["D21"])
    del d
    cfg = synth.MATCH_PARAMS
    out = {}

    def step():
        out["r"] = fm.match(X11, X21, D11, D21, cfg=cfg)

    ms = timed_steps(step, steps, warmup)
    px = B * N_PTS
    lib = _lib.load()
    F = D11.shape[-1]
    rays = torch.empty((B, H_IMG, W_IMG, 9), device=device)
    pts = torch.empty((B, N_PTS, 3), device=device)
    p0 = torch.empty((B, N_PTS, 2), device=device)
    conv = torch.ones((B, N_PTS), dtype=torch.bool, device=device)
    p1 = torch.empty((B, N_PTS, 2), dtype=torch.int64, device=device)
    v0 = torch.empty((B, N_PTS), dtype=torch.bool, device=device)
    Dh = torch.empty(D11.shape, dtype=torch.float16, device=device)
    sc = torch.rand((B, N_PTS), device=device).half()
    idx = torch.empty((B, N_PTS), dtype=torch.int64, device=device)
    stages = {
        "prep": (80, lambda: lib.frodo_match_prep(_ptr(X11), _ptr(X21), None, _ptr(rays), _ptr(pts),
                                                  _ptr(p0), B, H_IMG, W_IMG, _stream())),
        "occlusion": (50, lambda: lib.frodo_match_occlusion(
            _ptr(X11), _ptr(X21), _ptr(p0), _ptr(conv), 0.1, _ptr(p1), _ptr(v0), B, H_IMG, W_IMG,
            _stream())),
        "fp16_cast": (6 * F, lambda: lib.frodo_convert_f32_f16(_ptr(D11), _ptr(Dh), D11.numel(),
                                                               _stream())),
        "finish": (28, lambda: lib.frodo_match_finish(_ptr(p1), _ptr(sc), 1, _ptr(v0), 0.8,
                                                      _ptr(idx), _ptr(v0), B, N_PTS, W_IMG,
                                                      _stream())),
    }
    glue = {}
    glue_ms = 0.0
    for name, (bpp, fn) in stages.items():
        t = timed_steps(fn, 5, 2) / 5
        calls = 1  # the cast of D21 is fused into refine_matches' query load; only D11 is cast
        glue_ms += t * calls
        gbs = bpp * px / (t * 1e-3) / 1e9
        glue[name] = {"us": t * 1e3, "bytes_per_px": bpp, "achieved_gbs": gbs,
                      "roofline_frac": gbs / peak}
    return {"metric": "px-matches/s through match() (matching.py:8-10), fp32 inputs in HBM",
            "value": px * steps / (ms * 1e-3), "unit": "px/s", "pairs": B, "resolution": "512x384",
            "ms_per_step": ms / steps, "glue_share_of_call": glue_ms / (ms / steps),
            "glue_kernels": glue, "gpu_launches": 4 * steps,
            "kernels_per_call": "prep, iter_proj (+ occlusion test in its epilogue), fp16 cast of "
                                "D11, refine_matches (+ threshold and linearisation in its epilogue)"}


def cpu_baseline_gn(g_np, seconds=12.0):
    """Oracle port (C, OpenMP, all host threads) on a bounded sample of the same workload: the
    per-edge accumulation of a subset of the 256 edges (extrapolated to all of them) PLUS the
    assembly and the Cholesky solve of the full 441 x 441 system -- a whole iteration, as
    gn_kernels.cu:949-992 runs it."""
    from frodo_slam_b200 import synth
    from oracle import oracle
    gp = synth.GN_PARAMS
    cores = oracle.num_threads()
    ie, je, _ = oracle.edge_index(g_np["ii"], g_np["jj"])
    E = len(ie)
    ns = min(E, 2 * cores)
    t0 = time.perf_counter()
    passes = 0
    while True:
        Hs, gs = oracle.edge_blocks("rays", g_np["Twc"], g_np["Xs"], g_np["Cs"], ie[:ns], je[:ns],
                                    g_np["idx"][:ns], g_np["valid"][:ns], g_np["Q"][:ns],
                                    gp["sigma_ray"], gp["sigma_dist"], gp["C_conf"], gp["Q_conf"])
        passes += 1
        if time.perf_counter() - t0 > seconds:
            break
    dt = (time.perf_counter() - t0) / passes
    # assembly + solve of the FULL system (the sampled blocks tiled over all edges: timing only)
    rep = (E + ns - 1) // ns
    Hs_all = np.tile(Hs, (1, rep, 1, 1))[:, :E]
    gs_all = np.tile(gs, (1, rep, 1))[:, :E]
    K = g_np["Xs"].shape[0]
    t1 = time.perf_counter()
    H, gvec = oracle.assemble(Hs_all, gs_all, ie, je, K, gp["num_fix"])
    H = H + 1e-3 * np.abs(H).max() * np.eye(H.shape[0])  # tiled blocks: keep it factorisable
    oracle.chol_solve(H, gvec)
    t_solve = time.perf_counter() - t1
    t_iter = dt * E / ns + t_solve
    return {"value": 1.0 / t_iter, "unit": "iters/s", "cores": cores, "kind": "port",
            "sample": f"{ns} of {E} edges x {N_PTS} points, {passes} passes (accumulation "
                      f"extrapolated x{E / ns:.0f}: {dt * E / ns:.2f} s) + assembly and Cholesky of "
                      f"the full {H.shape[0]}x{H.shape[0]} system ({t_solve * 1e3:.1f} ms)"}


def cpu_baseline_gn_torch(g_np, edges=4):
    """The PyTorch-CPU formulation (oracle/torch_cpu.py, SURVEY 8(d) "C1") of one whole
    gauss_newton_rays iteration on all host cores: accumulation of `edges` of the 256 edges
    (extrapolated), assembly and Cholesky solve of the full system."""
    from frodo_slam_b200 import synth
    from oracle import torch_cpu
    gp = synth.GN_PARAMS
    cores = torch_cpu.use_all_cores()
    t = {k: torch.from_numpy(np.ascontiguousarray(v)) for k, v in g_np.items()}
    E = t["ii"].shape[0]
    valid = t["valid"].bool()
    args = (t["Twc"], t["Xs"], t["Cs"], t["ii"], t["jj"], t["idx"], valid, t["Q"], gp["sigma_ray"],
            gp["sigma_dist"], gp["C_conf"], gp["Q_conf"], gp["num_fix"])
    sel = list(range(0, E, max(1, E // edges)))[:edges]
    torch_cpu.gn_rays_system(*args, edges=sel[:1])  # warm-up
    t0 = time.perf_counter()
    H, g = torch_cpu.gn_rays_system(*args, edges=sel)
    t_acc = (time.perf_counter() - t0) * E / len(sel)
    H = H + 1e-3 * H.abs().max() * torch.eye(H.shape[0], dtype=H.dtype)
    t1 = time.perf_counter()
    torch_cpu.gn_solve(H, g)
    t_solve = time.perf_counter() - t1
    return {"value": 1.0 / (t_acc + t_solve), "unit": "iters/s", "cores": cores,
            "kind": "port (PyTorch-CPU formulation)",
            "sample": f"{len(sel)} of {E} edges x {N_PTS} points (accumulation extrapolated: "
                      f"{t_acc:.2f} s) + Cholesky solve ({t_solve * 1e3:.1f} ms)"}


def cpu_baseline_matching():
    """BASELINE.json configs[0]: iter_proj + refine_matches on ONE 512x384 pair on the host
    cores, C/OpenMP oracle port and PyTorch-CPU formulation."""
    from frodo_slam_b200 import synth
    from oracle import oracle, torch_cpu
    mp = synth.MATCH_PARAMS
    d = synth.matching_pairs(1, H_IMG, W_IMG, seed=0)
    rays, pts, p0 = synth.prep_for_iter_proj(d["X11"], d["X21"])
    D11 = d["D11"].half().contiguous()
    D21 = d["D21"].half().reshape(1, N_PTS, -1).contiguous()
    out = {}
    t0 = time.perf_counter()
    p, _ = oracle.iter_proj(rays.numpy(), pts.numpy(), p0.numpy(), mp["max_iter"],
                            mp["lambda_init"], mp["convergence_thresh"])
    t1 = time.perf_counter()
    p1 = torch.from_numpy(p).long()
    oracle.refine_matches(D11.numpy(), D21.numpy(), p1.numpy(), mp["radius"], mp["dilation_max"])
    t2 = time.perf_counter()
    out["c_port"] = {"value": N_PTS / (t2 - t0), "unit": "px-matches/s", "cores": oracle.num_threads(),
                     "kind": "port", "iter_proj_s": t1 - t0, "refine_matches_s": t2 - t1,
                     "sample": "one 512x384 pair (configs[0]), whole ops"}
    cores = torch_cpu.use_all_cores()
    t0 = time.perf_counter()
    pt, _ = torch_cpu.iter_proj(rays, pts, p0, mp["max_iter"], mp["lambda_init"],
                                mp["convergence_thresh"])
    t1 = time.perf_counter()
    torch_cpu.refine_matches(D11, D21, pt.long(), mp["radius"], mp["dilation_max"])
    t2 = time.perf_counter()
    out["torch_cpu"] = {"value": N_PTS / (t2 - t0), "unit": "px-matches/s", "cores": cores,
                        "kind": "port (PyTorch-CPU formulation)", "iter_proj_s": t1 - t0,
                        "refine_matches_s": t2 - t1,
                        "sample": "one 512x384 pair (configs[0]), whole ops"}
    return out


def run_other_configs(dev, be, iters, quiet=False):
    """The smaller graphs of BASELINE.json: configs[1] (8-keyframe sliding window, E = 24,
    gauss_newton_rays) and configs[3] (32 keyframes, E = 128, gauss_newton_calib, TUM-like
    intrinsics) at 512x384, through the backend module `be` (ours or the reference's): iters/s of
    a 10-iteration call with delta_thresh = 0, inputs resident on the device."""
    from frodo_slam_b200 import synth
    gp = synth.GN_PARAMS
    out = {}
    for name, cfgname, mode in (("window8_rays", "window8", "rays"), ("calib32", "calib32", "calib")):
        K, und = synth.graph_edges(cfgname)
        g = synth.keyframe_graph(K, H_IMG, W_IMG, und, seed=1, device=dev)
        T0 = g["Twc"].clone()

        def call():
            T = T0.clone()
            if mode == "rays":
                be.gauss_newton_rays(T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"],
                                     g["valid_match"], g["Q"], gp["sigma_ray"], gp["sigma_dist"],
                                     gp["C_conf"], gp["Q_conf"], gp["num_fix"], iters, 0.0)
            else:
                be.gauss_newton_calib(T, g["Xs"], g["Cs"], g["K"], g["ii"], g["jj"],
                                      g["idx_ii2jj"], g["valid_match"], g["Q"], H_IMG, W_IMG,
                                      gp["pixel_border"], gp["depth_eps"], gp["sigma_pixel"],
                                      gp["sigma_depth"], gp["C_conf"], gp["Q_conf"], iters, 0.0)

        n = 5
        if quiet:
            with _QuietStdout():
                ms = timed_steps(call, n, 2)
        else:
            ms = timed_steps(call, n, 2)
        E = int(g["ii"].shape[0])
        out[name] = {"keyframes": K, "directed_edges": E, "iters_per_s": n * iters / (ms * 1e-3),
                     "ms_per_call": ms / n,
                     "algorithmic_gb_per_iteration": 45 * N_PTS * E / 1e9}
        del g
        torch.cuda.empty_cache()
    return out


def run_tracking(dev):
    """The tracker's frame-to-keyframe pose solve (tracker.py:397-517) on one
    512x384 frame through frodo_slam_b200.tracking (one persistent cooperative kernel): time per
    solve with the reference's stopping rule, and per iteration from fixed 1- and 50-iteration
    runs.  profiles/bench_track.py adds the PyTorch formulation of the same loop."""
    from frodo_slam_b200 import synth, tracking
    c = synth.tracking_case(H_IMG, W_IMG, seed=7, pose_noise=0.01)
    t = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in c.items()}
    tp = dict(synth.TRACK_PARAMS)
    fixed = dict(tp, rel_error=0.0, delta_norm=0.0)

    def call(cfg):
        return tracking.track_pose("rays", t["Xf_all"], t["Xk"], t["T_WCf"], t["T_WCk"], t["Qk"],
                                   t["valid"], idx_f2k=t["idx_f2k"], cfg=cfg)

    iters = int(call(tp)["info"][0].item())
    n = 50
    us = 1e3 * timed_steps(lambda: call(tp), n, 5) / n
    us1 = 1e3 * timed_steps(lambda: call(dict(fixed, max_iters=1)), n, 3) / n
    us50 = 1e3 * timed_steps(lambda: call(fixed), n, 3) / n
    return {"metric": "tracker pose solves/s (opt_pose_ray_dist_sim3), 512x384 frame",
            "value": 1e6 / us, "unit": "solves/s", "us_per_solve": us, "iterations": iters,
            "us_per_iteration": (us50 - us1) / 49.0, "points": N_PTS,
            "prologue_bytes": N_PTS * 37, "gpu_launches_per_solve": 1,
            "note": "after a one-off 37 B/point prologue the points stay in shared memory; "
                    "per-iteration global traffic is 36 doubles per CTA"}


def run_arena_e2e(dev, g, Twc0, idx, valid, Q, gp, iters, steps):
    """The same solve fed the way a SLAM session feeds it: the graph lives in
    a device-resident KeyframeGraphArena; each step the NEWEST keyframe arrives from pinned
    host memory with its edges (pointmap 3.1 MB + 2.56 MB per directed edge), replaces its
    previous version in the arena, the whole 64-keyframe graph is solved and the poses are read
    back.  The reference re-stacks and re-uploads everything here (global_opt.py:186-224)."""
    from frodo_slam_b200.graph import KeyframeGraphArena
    K = g["Xs"].shape[0]
    E = g["ii"].shape[0]
    nu = E // 2
    ii, jj = g["ii"].tolist(), g["jj"].tolist()
    A = KeyframeGraphArena(N_PTS, K, nu, device=dev)
    for k in range(K):
        A.put_keyframe(k, g["Xs"][k], g["Cs"][k], Twc0[k])
    A.add_edges(ii[:nu], jj[:nu], idx[:nu], idx[nu:], valid[:nu], valid[nu:], Q[:nu], Q[nu:])
    new = K - 1
    sel = [e for e in range(nu) if ii[e] == new or jj[e] == new]
    selb = [nu + e for e in sel]
    pin = lambda t: t.cpu().pin_memory()  # noqa: E731
    host = dict(X=pin(g["Xs"][new]), C=pin(g["Cs"][new]), T=pin(Twc0), idx_f=pin(idx[sel]),
                idx_b=pin(idx[selb]), v_f=pin(valid[sel]), v_b=pin(valid[selb]), Q_f=pin(Q[sel]),
                Q_b=pin(Q[selb]))
    out_host = torch.empty(K, 8).pin_memory()
    e_ii, e_jj = [ii[e] for e in sel], [jj[e] for e in sel]
    h2d = sum(v.numel() * v.element_size() for v in host.values())
    ids = list(range(K))

    def step():
        d = {k: v.to(dev, non_blocking=True) for k, v in host.items()}
        A.Twc.copy_(d["T"])  # restart from the same perturbed poses (2 KB)
        A.remove_keyframe(new)
        A.put_keyframe(new, d["X"], d["C"], d["T"][new])
        A.add_edges(e_ii, e_jj, d["idx_f"], d["idx_b"], d["v_f"], d["v_b"], d["Q_f"], d["Q_b"])
        A.solve_rays(gp["sigma_ray"], gp["sigma_dist"], gp["C_conf"], gp["Q_conf"], gp["num_fix"],
                     iters, 0.0)
        out_host.copy_(A.poses(ids), non_blocking=True)

    ms = timed_steps(step, steps, 2)
    del A
    return {"value": steps * iters / (ms * 1e-3), "unit": "iters/s", "ms_per_step": ms / steps,
            "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": out_host.numel() * 4,
            "new_edges_per_step": len(sel),
            "api": "frodo_slam_b200.graph.KeyframeGraphArena (put_keyframe + add_edges + solve_rays)",
            "note": "graph resident on the device; only the newest keyframe and its edges cross "
                    "PCIe each step"}


def main_ours(args):
    import torch.distributed as dist
    import mast3r_slam_backends as be
    from frodo_slam_b200 import _lib, dist as fdist, synth
    rank, world, local = dist_env()
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    barrier = (lambda: dist.barrier()) if world > 1 else None
    peak, peak_src = measured_peak_gbs()
    prm, gp = gn_params()
    K, g = build_graph(dev)
    E = g["ii"].shape[0]
    lo, hi = fdist.edge_slice(E, rank, world)
    idx = g["idx_ii2jj"].view(E, N_PTS)
    valid = g["valid_match"].view(E, N_PTS)
    Q = g["Q"].view(E, N_PTS)
    Twc0 = g["Twc"].clone()
    Twc = g["Twc"].clone()
    sg = fdist.ShardedGaussNewton(prm, Twc, g["Xs"], g["Cs"], None, g["ii"], g["jj"],
                                  idx[lo:hi].contiguous(), valid[lo:hi].contiguous(),
                                  Q[lo:hi].contiguous(), lo, hi, gp["num_fix"])
    iters = args.gn_iters
    exchange = ("none (single GPU)" if world == 1 else
                "per-edge blocks pushed to every rank from the accumulate kernel (NVLink peer "
                "memory, 280 B per edge per rank), one C call per solve" if sg.blocks is not None else
                "dense [H|g] all-reduce: peer-memory kernel" if sg.peer is not None else
                "dense [H|g] all-reduce: NCCL")

    def step():
        Twc.copy_(Twc0)
        sg.solve(iters, 0.0)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms = timed_steps(step, args.steps, args.warmup, barrier,
                     on_start=sampler.mark if rank == 0 else None)
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    clocks = sampler.stop() if rank == 0 else None
    assert int(sg.flags[1].item()) == iters and int(sg.flags[2].item()) == 0, "GN did not run clean"
    value = args.steps * iters / (ms * 1e-3)

    # Parity of THIS run's code path (every N): the poses one step produces, against the
    # single-GPU drop-in operator on the full graph, against the unmodified reference backend
    # (oracle/_ref, gn_kernels.cu:909-995 + Eigen solve) when it is on the box, and across ranks.
    def _rel(a, b):
        return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-300))

    step()
    T_run = Twc.clone()
    T_single = Twc0.clone()
    be.gauss_newton_rays(T_single, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"],
                         g["valid_match"], g["Q"], gp["sigma_ray"], gp["sigma_dist"], gp["C_conf"],
                         gp["Q_conf"], gp["num_fix"], iters, 0.0)
    parity = {"pose_rel_err_vs_single_gpu_op": _rel(T_run, T_single),
              "pose_rel_err_vs_reference": None, "ranks_bit_identical": True, "tolerance": 1e-4}
    if world > 1:
        allT = [torch.empty_like(T_run) for _ in range(world)]
        dist.all_gather(allT, T_run)
        parity["ranks_bit_identical"] = bool(all(torch.equal(allT[0], x) for x in allT))
    if rank == 0:
        ref_mod = load_reference_backend()
        if ref_mod is not None:
            T_ref = Twc0.clone()
            with _QuietStdout():
                ref_mod.gauss_newton_rays(T_ref, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"],
                                          g["valid_match"], g["Q"], gp["sigma_ray"],
                                          gp["sigma_dist"], gp["C_conf"], gp["Q_conf"],
                                          gp["num_fix"], iters, 0.0)
                torch.cuda.synchronize()
            parity["pose_rel_err_vs_reference"] = _rel(T_run, T_ref)
            parity["pose_moved_rel"] = _rel(Twc0, T_ref)  # how far the solve moved the poses
            assert parity["pose_rel_err_vs_reference"] <= 1e-4, parity
    assert parity["pose_rel_err_vs_single_gpu_op"] <= 1e-4 and parity["ranks_bit_identical"], parity

    # dominant kernel alone (accumulate, no assembly): CUDA events on the launch stream
    lib = _lib.load()
    import ctypes
    from frodo_slam_b200.backends import _ptr, _stream
    Twc.copy_(Twc0)

    def acc_only():
        _lib.check(lib.frodo_gn_accumulate(
            ctypes.byref(prm), _ptr(Twc), _ptr(g["Xs"]), _ptr(g["Cs"]), None, sg.xc, _ptr(sg.ii_row),
            _ptr(sg.jj_row), _ptr(sg.idx), _ptr(sg.valid), _ptr(sg.Q), K, N_PTS, sg.E_local,
            gp["num_fix"], None, None, None, _ptr(sg.ws), sg.nbytes, _stream()))

    n_acc = max(args.steps * iters, 10)
    ms_acc = timed_steps(acc_only, n_acc, 3) / n_acc
    bytes_launch = GN_BYTES_PER_POINT * N_PTS * sg.E_local
    achieved = bytes_launch / (ms_acc * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": "gn_accumulate_packed_kernel<rays>", "achieved": achieved,
                "peak": peak, "peak_source": peak_src, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None, "us_per_launch": ms_acc * 1e3,
                "algorithmic_bytes_per_launch": bytes_launch,
                "share_of_step": ms_acc * iters / (ms / args.steps)}
    prof = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(prof):
        try:
            full = json.load(open(prof)).get("gn_accumulate_rays_bytes_per_launch")
            # measured on the full 256-edge launch; a rank's launch covers E_local edges
            roofline["traffic"] = None if full is None else int(full * (hi - lo) / E)
        except Exception:
            pass

    # end to end: host (pinned) buffers -> operator -> poses back on the host
    host = {k: g[k].cpu().pin_memory() for k in ("Xs", "Cs", "ii", "jj")}
    host["Twc"] = Twc0.cpu().pin_memory()
    host["idx"] = idx[lo:hi].cpu().pin_memory()
    host["valid"] = valid[lo:hi].cpu().pin_memory()
    host["Q"] = Q[lo:hi].cpu().pin_memory()
    out_host = torch.empty_like(host["Twc"]).pin_memory()
    h2d = sum(v.numel() * v.element_size() for v in host.values())
    d2h = out_host.numel() * out_host.element_size()
    idx_full, valid_full, Q_full = (idx, valid, Q) if world == 1 else (None, None, None)
    del sg, idx, valid, Q
    g_full_E = E

    def e2e_step():
        d = {k: v.to(dev, non_blocking=True) for k, v in host.items()}
        if world == 1:
            be.gauss_newton_rays(d["Twc"], d["Xs"], d["Cs"], d["ii"], d["jj"], d["idx"],
                                 d["valid"].view(-1, N_PTS, 1), d["Q"], gp["sigma_ray"],
                                 gp["sigma_dist"], gp["C_conf"], gp["Q_conf"], gp["num_fix"],
                                 iters, 0.0)
        else:
            s = fdist.ShardedGaussNewton(prm, d["Twc"], d["Xs"], d["Cs"], None, d["ii"], d["jj"],
                                         d["idx"], d["valid"], d["Q"], lo, hi, gp["num_fix"])
            s.solve(iters, 0.0)
        out_host.copy_(d["Twc"], non_blocking=True)

    e2e_steps = max(2, min(args.steps, 5))
    ms_serial = timed_steps(e2e_step, e2e_steps, 1, barrier)

    # The same call with the next step's host->device copy overlapped with this step's solve
    # (two device buffer sets, one copy stream): every step still copies all of its inputs from
    # pinned host memory and reads its poses back, inside the timed region.
    copy_stream = torch.cuda.Stream(device=dev)
    main_stream = torch.cuda.current_stream(dev)
    sets = [{k: torch.empty_like(v, device=dev) for k, v in host.items()} for _ in range(2)]
    copied = [torch.cuda.Event() for _ in range(2)]
    freed = [torch.cuda.Event() for _ in range(2)]

    def upload(b):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(freed[b])
            for k, v in host.items():
                sets[b][k].copy_(v, non_blocking=True)
            copied[b].record(copy_stream)

    def solve_set(b):
        d = sets[b]
        main_stream.wait_event(copied[b])
        if world == 1:
            be.gauss_newton_rays(d["Twc"], d["Xs"], d["Cs"], d["ii"], d["jj"], d["idx"],
                                 d["valid"].view(-1, N_PTS, 1), d["Q"], gp["sigma_ray"],
                                 gp["sigma_dist"], gp["C_conf"], gp["Q_conf"], gp["num_fix"],
                                 iters, 0.0)
        else:
            sh = fdist.ShardedGaussNewton(prm, d["Twc"], d["Xs"], d["Cs"], None, d["ii"], d["jj"],
                                          d["idx"], d["valid"], d["Q"], lo, hi, gp["num_fix"])
            sh.solve(iters, 0.0)
        out_host.copy_(d["Twc"], non_blocking=True)
        freed[b].record(main_stream)

    def e2e_pipelined(n):
        torch.cuda.synchronize()
        if barrier:
            barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(main_stream)
        copy_stream.wait_event(e0)
        for b in range(2):
            freed[b].record(main_stream)
        upload(0)
        for k in range(n):
            if k + 1 < n:
                upload((k + 1) % 2)
            solve_set(k % 2)
        e1.record(main_stream)
        torch.cuda.synchronize()
        if barrier:
            barrier()
        return e0.elapsed_time(e1)

    pipe_steps = max(5, args.steps)
    e2e_pipelined(2)
    ms_e2e = e2e_pipelined(pipe_steps)
    t = torch.tensor([ms_e2e, ms_serial], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_e2e, ms_serial = float(t[0].item()), float(t[1].item())
    del sets
    e2e = {"value": pipe_steps * iters / (ms_e2e * 1e-3), "unit": "iters/s",
           "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e / pipe_steps,
           "steps": pipe_steps,
           "h2d_gbs": h2d / (ms_e2e / pipe_steps * 1e-3) / 1e9,  # PCIe rate the step sustains
           "bound": (f"PCIe host->device copy ({h2d / 1e6:.0f} MB per step at "
                     f"{h2d / (ms_e2e / pipe_steps * 1e-3) / 1e9:.0f} GB/s; the solve, "
                     f"{ms / args.steps:.1f} ms, hides under the upload)"),
           "overlap": "step n+1's host->device copy (copy stream, second buffer set) runs under "
                      "step n's solve; every step copies all inputs and reads its poses back",
           "serial_value": e2e_steps * iters / (ms_serial * 1e-3),
           "serial_ms_per_step": ms_serial / e2e_steps,
           "api": "mast3r_slam_backends.gauss_newton_rays" if world == 1
                  else "frodo_slam_b200.dist.ShardedGaussNewton"}

    if world == 1:
        e2e["arena"] = run_arena_e2e(dev, g, Twc0, idx_full, valid_full, Q_full, gp, iters,
                                     max(5, args.steps))
    cpu = None
    g_np = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        g_np = {"Twc": Twc0.cpu().numpy(), "Xs": g["Xs"].cpu().numpy(), "Cs": g["Cs"].cpu().numpy(),
                "ii": g["ii"].cpu().numpy(), "jj": g["jj"].cpu().numpy(),
                "idx": host["idx"].numpy(), "valid": host["valid"].numpy().astype(np.uint8),
                "Q": host["Q"].numpy()}
    del g, host
    torch.cuda.empty_cache()
    matching = None
    if not args.no_matching:
        Bm = max(1, args.pairs // world)
        matching = run_matching(dev, Bm, max(2, min(args.steps, 5)), 3, peak)
        if world > 1:
            tv = torch.tensor([matching["value"]], device=dev, dtype=torch.float64)
            dist.all_reduce(tv, op=dist.ReduceOp.SUM)
            matching["value"] = float(tv.item())
            matching["pairs"] = Bm * world
            matching["note"] = "batch split over ranks, no collective (replicas only)"
        if world == 1:
            torch.cuda.empty_cache()
            matching["match_call"] = run_match_pipeline(dev, min(Bm, 64), 3, 2, peak)
            torch.cuda.empty_cache()
            # the other resolution of BASELINE.json configs[4] (batched tracking sweep)
            m224 = run_matching(dev, Bm, 3, 2, peak, hw=(224, 224))
            matching["sweep_224x224"] = {k: m224[k] for k in ("value", "unit", "pairs", "resolution",
                                                              "ms_per_step", "iter_proj",
                                                              "refine_matches")}
    cpu_torch = None
    if g_np is not None:
        cpu = cpu_baseline_gn(g_np)
        cpu_torch = cpu_baseline_gn_torch(g_np)
        if matching is not None:
            matching["cpu_baseline"] = cpu_baseline_matching()
    tracking = run_tracking(dev) if (world == 1 and not args.no_matching) else None
    other = run_other_configs(dev, be, iters) if (world == 1 and not args.no_matching) else None

    if rank == 0:
        line = {
            "metric": "GN-rays iters/s, 64-KF 512x384", "value": value, "unit": "iters/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32 (fp64 assembly/solve)", "data": "synthetic",
            "config": {"workload": "configs[2]: 64-keyframe graph, 63 chain + 65 loop-closure edges "
                                   "both directions (E=256), 512x384 pointmaps, gauss_newton_rays, "
                                   f"{iters} iterations per step, delta_thresh=0",
                       "edges_per_gpu": hi - lo, "sharding": f"edges/{world}", "exchange": exchange,
                       "l2": "inputs larger than L2 (2.27 GB streamed per iteration)"},
            "roofline": roofline, "cpu_baseline": cpu, "cpu_baseline_torch": cpu_torch,
            "e2e": e2e, "clocks": clocks,
            "parity": parity,
            "gpu_launches": args.steps * (1 + 2 * iters),  # pack + (accumulate, solve) per iteration
            "matching": matching,
            "tracking": tracking,
            "other_configs": other,
        }
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


class _QuietStdout:
    """The reference prints its solver log with std::cout every iteration: silence fd 1."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        self.null = os.open(os.devnull, os.O_WRONLY)
        os.dup2(self.null, 1)

    def __exit__(self, *a):
        os.dup2(self.saved, 1)
        os.close(self.null)
        os.close(self.saved)


def load_reference_backend():
    import glob
    import importlib.util
    so = glob.glob(os.path.join(ROOT, "oracle", "_ref", "mast3r_slam_backends*.so"))
    if not so:
        return None
    try:
        spec = importlib.util.spec_from_file_location("mast3r_slam_backends", so[0])
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod
    except Exception:
        return None


def main_reference(args):
    """Reference arm: the UNMODIFIED reference backend (oracle/_ref, its own pybind API and
    stock CUDA + Eigen code path) on the same graph when it is built and a GPU is present;
    otherwise the oracle port on the host cores.  Rank 0 only."""
    rank, world, local = dist_env()
    if rank != 0:
        return
    from frodo_slam_b200 import synth
    gp = synth.GN_PARAMS
    iters = args.gn_iters
    ref = load_reference_backend() if torch.cuda.is_available() else None
    dev = torch.device("cuda", local) if torch.cuda.is_available() else torch.device("cpu")
    if torch.cuda.is_available():
        torch.cuda.set_device(local)
    K, g = build_graph(dev)
    E = g["ii"].shape[0]
    g_np = {"Twc": g["Twc"].cpu().numpy(), "Xs": g["Xs"].cpu().numpy(), "Cs": g["Cs"].cpu().numpy(),
            "ii": g["ii"].cpu().numpy(), "jj": g["jj"].cpu().numpy(),
            "idx": g["idx_ii2jj"].view(E, N_PTS).cpu().numpy(),
            "valid": g["valid_match"].view(E, N_PTS).cpu().numpy().astype(np.uint8),
            "Q": g["Q"].view(E, N_PTS).cpu().numpy()}
    cfg = {"workload": "configs[2]: 64-keyframe graph, 63 chain + 65 loop-closure edges both "
                       f"directions (E=256), 512x384 pointmaps, gauss_newton_rays, {iters} "
                       "iterations per step, delta_thresh=0"}
    if ref is not None:
        Twc0 = g["Twc"].clone()

        def step():
            T = Twc0.clone()
            ref.gauss_newton_rays(T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"],
                                  g["valid_match"], g["Q"], gp["sigma_ray"], gp["sigma_dist"],
                                  gp["C_conf"], gp["Q_conf"], gp["num_fix"], iters, 0.0)

        with _QuietStdout():
            ms = timed_steps(step, args.steps, max(1, min(args.warmup, 2)))
        value = args.steps * iters / (ms * 1e-3)
        kind, cores = "reference", os.cpu_count()
        sample = ("full workload on the same GPU through the reference's own CUDA kernels + host "
                  "Eigen solve (the reference has no CPU implementation of this path)")
        cfg["device"] = torch.cuda.get_device_name(local)
        steps_done = args.steps
    else:
        del g
        cb = cpu_baseline_gn(g_np, seconds=10.0 * max(1, args.steps))
        value, kind, cores, sample = cb["value"], cb["kind"], cb["cores"], cb["sample"]
        ms, steps_done = 1e3 * iters / value, 1
    line = {"impl": "reference", "metric": "GN-rays iters/s, 64-KF 512x384", "value": value,
            "unit": "iters/s", "n_gpus": 1, "steps": steps_done, "warmup": args.warmup,
            "ms_per_step": ms / steps_done, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32 (fp64 assembly/solve)", "data": "synthetic",
            "config": cfg,
            "cpu_baseline": {"value": value, "unit": "iters/s", "cores": cores, "kind": kind,
                             "sample": sample},
            "e2e": {"value": value, "unit": "iters/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0}}
    if ref is not None and not args.no_cpu_baseline:
        line["cpu_port"] = cpu_baseline_gn(g_np, seconds=10.0)
    if ref is not None and not args.no_matching:
        # the other half of BASELINE.json's metric, through the reference's own kernels
        del g
        torch.cuda.empty_cache()
        m = run_matching(dev, args.pairs, max(2, min(args.steps, 5)), 2, measured_peak_gbs()[0],
                         be=ref)
        m["gpu_launches"] = None
        line["matching"] = m
        line["other_configs"] = run_other_configs(dev, ref, iters, quiet=True)
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--gn-iters", type=int, default=10)
    ap.add_argument("--pairs", type=int, default=256, help="frame pairs of the matching sweep")
    ap.add_argument("--no-matching", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    # stdout carries exactly ONE line, the JSON: anything libraries write to fd 1 meanwhile
    # (NCCL's version banner, the reference's solver log) is routed to stderr.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    out = io.StringIO()
    with contextlib.redirect_stdout(out):
        if args.impl == "reference":
            main_reference(args)
        else:
            main_ours(args)
    sys.stdout.flush()
    lines = [ln for ln in out.getvalue().splitlines() if ln.strip()]
    json_lines = [ln for ln in lines if ln.lstrip().startswith("{")]
    for ln in lines:
        if ln not in json_lines:
            print(ln, file=sys.stderr)
    if json_lines:
        os.write(real_stdout, (json_lines[-1] + "\n").encode())
    os.close(real_stdout)


if __name__ == "__main__":
    main()
