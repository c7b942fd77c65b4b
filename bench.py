#!/usr/bin/env python3
"""bench.py -- KITTI-shape ORB extraction throughput on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one pass of the whole extraction path over a batch of synthetic KITTI-shape frames
(1241x376, 2000 keypoints, 8 levels, scale 1.2, FAST 20/7 = Examples/Stereo/KITTI00-02.yaml of
the reference).  Frames are independent, so with N GPUs every rank processes its own shard of the
sequence and nothing is exchanged ("scaling": "weak").

  value   frames/s with the frames already resident in HBM, timed with CUDA events on the
          library's own stream (borb_timer_*), max over ranks
  e2e     frames/s through the host-facing C ABI (borb_submit_host / borb_wait) with page-locked
          host buffers: H2D of every frame and D2H of every keypoint/descriptor inside the timing
  roofline  the slowest kernel stage against the measured HBM peak (MEASURED_PEAKS.json)
  cpu_baseline  the reference's own ORBextractor.cc (oracle/_ref, compiled from the reference tree
          against the OpenCV-4.13-pinned primitives) on this host's cores, bounded sample

`--impl reference` times that CPU implementation alone, same metric and config.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "KITTI-shape ORB extract frames/s"
NFEATURES, SCALE, NLEVELS, INI_TH, MIN_TH = 2000, 1.2, 8, 20, 7
ROWS, COLS = 376, 1241
UNIQUE = 16  # distinct synthetic frames per rank; a step's batch cycles through them


def level_pixels() -> int:
    s, tot = 1.0, 0
    f32 = np.float32
    sc = f32(1.0)
    for l in range(NLEVELS):
        inv = f32(1.0) / sc
        tot += int(np.rint(f32(COLS) * inv)) * int(np.rint(f32(ROWS) * inv))
        sc = f32(float(sc) * float(f32(SCALE)))
    return tot


def algorithmic_bytes(n_kp: float) -> float:
    """read the frame once, write the 8 un-padded levels, 60 B per keypoint."""
    return COLS * ROWS + level_pixels() + 60.0 * n_kp


def measured_peak():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock and throttle reasons sampled WHILE the timed region runs: NVML every 5 ms from a
    thread (a timed region is tens of milliseconds; an nvidia-smi subprocess cannot sample that
    fast, least of all on an 8-GPU box); nvidia-smi -lms as the fallback when pynvml is missing."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    BITS = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}

    def __init__(self, gpu: int):
        self.gpu, self.rows, self.proc = gpu, [], None
        self.sm, self.mx, self.reasons, self.stop_flag, self.thread, self.src = [], [], set(), False, None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            # LOCAL_RANK indexes CUDA_VISIBLE_DEVICES; NVML indexes physical devices
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = int(vis.split(",")[gpu]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else gpu
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.src = "nvml"
        except Exception:
            self.nv = None

    def _poll(self):
        nv, h = self.nv, self.h
        while not self.stop_flag:
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                self.mx.append(float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)))
                get = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
                mask = int(get(h))
                for name, bit in self.BITS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.005)

    def start(self):
        if self.nv is not None:
            self.stop_flag = False
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
            self.src = "nvidia-smi"
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.thread is not None:
            self.stop_flag = True
            self.thread.join()
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
        sm, mx, reasons = list(self.sm), list(self.mx), set(self.reasons)
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": self.src}


def bind_near_gpu(local: int) -> str:
    """Run this rank (and so first-touch its pinned buffers) on the CPU cores of the NUMA node the
    GPU hangs off: with several GPUs per box the host<->device copies of the e2e measurement
    otherwise cross the socket interconnect.  Returns a note for the JSON line."""
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = int(vis.split(",")[local]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else local
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(idx)).busId
        bus = (bus.decode() if isinstance(bus, bytes) else bus).lower()
        if len(bus.split(":")[0]) == 8:
            bus = bus[4:]
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        if node < 0:
            return "numa: single node"
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return f"numa: bound to node {node} ({len(cpus)} cpus)"
        return f"numa: node {node} has no allowed cpus"
    except Exception as e:
        return f"numa: not bound ({type(e).__name__})"


def dist_setup(n_gpus: int):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, world, local


def all_max(x: float, world: int) -> float:
    if world == 1:
        return x
    import torch
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def all_sum(x: float, world: int) -> float:
    if world == 1:
        return x
    import torch
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def barrier(world: int):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()


# ----------------------------------------------------------------------------- CPU reference arm

def cpu_reference_run(procs: int, frames_per_proc: int) -> dict:
    """Time the reference's CPU ORBextractor on this host: `procs` OS processes, one extractor and
    one frame at a time each (oracle/cpu_bench.py, run as a child so it never shares an address
    space with CUDA).  Returns its JSON: frames_per_s, frames, kind ("reference" = the reference's
    own ORBextractor.cc from oracle/_ref, "port" = the restatement), ms_per_frame_per_core."""
    out = subprocess.run([sys.executable, "-m", "oracle.cpu_bench", "--procs", str(procs),
                          "--frames-per-proc", str(frames_per_proc), "--rows", str(ROWS), "--cols", str(COLS),
                          "--nfeatures", str(NFEATURES)], cwd=str(ROOT), capture_output=True, text=True, check=True)
    return json.loads(out.stdout.strip().splitlines()[-1])


CPU_FRAMES_PER_PROC = 32


def cpu_baseline_stable(cores: int, runs: int = 5) -> dict:
    """All-cores reference throughput: every worker pinned to its own core (oracle/cpu_bench.py), 32 frames per
    process, median of `runs` runs with min and max reported."""
    rs = [cpu_reference_run(cores, CPU_FRAMES_PER_PROC) for _ in range(runs)]
    v = sorted(r["frames_per_s"] for r in rs)
    return {"value": v[len(v) // 2], "unit": "frames/s", "cores": cores, "kind": rs[0]["kind"],
            "min": v[0], "max": v[-1], "spread": (v[-1] - v[0]) / v[len(v) // 2],
            "sample": f"{rs[0]['frames']} KITTI-shape frames per run ({cores} pinned processes x {CPU_FRAMES_PER_PROC} frames, one extractor "
                      f"each), median of {runs} runs, min {v[0]:.0f} max {v[-1]:.0f}"}


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    cores = len(os.sched_getaffinity(0))
    per_proc = CPU_FRAMES_PER_PROC
    for _ in range(min(args.warmup, 1)):
        cpu_reference_run(cores, 1)
    tot, wall, kind = 0, 0.0, "reference"
    for _ in range(args.steps):
        r = cpu_reference_run(cores, per_proc)
        tot += r["frames"]
        wall += r["wall_s"]
        kind = r["kind"]
    value = tot / wall  # = mean over the steps; the steps' own rates are in cpu_baseline.step_frames_per_s
    per_step = cores * per_proc
    sample = (f"bounded sample of the workload: {tot} frames ({per_step} per step), {cores} processes, "
              "one extractor per process, one frame at a time each")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": wall / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic", "config": bench_config(args.frames_per_step),
            "cpu_baseline": {"value": value, "unit": "frames/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def bench_config(frames_per_step: int) -> dict:
    return {"workload": "K2k: KITTI mono 1241x376, 2000 kp, 8 levels, scale 1.2, FAST 20/7 "
                        "(Examples/Stereo/KITTI00-02.yaml)",
            "frames_per_step": frames_per_step, "cols": COLS, "rows": ROWS, "nfeatures": NFEATURES,
            "l2": "inputs larger than L2: one step reads frames_per_step x 466 KB and streams "
                  "~8 MB of per-frame intermediates through HBM",
            "sharding": "frame-wise, no collective"}


# ----------------------------------------------------------------------------- GPU arm

STAGE_KERNEL = {"pyramid": "k_pyr_level0 + 7 x k_pyr_resize", "fast": "k_fast_cells", "quadtree": "k_quadtree",
                "blur": "k_blur_strips", "describe": "k_orient_describe"}


def stage_algorithmic_bytes(n_kp: float, n_cand: float) -> dict:
    """Algorithmic bytes per FRAME of each stage: every stage reads its input
    and writes its output once."""
    lp = level_pixels()
    return {
        "pyramid": COLS * ROWS + lp,            # read the frame, write every level once
        "fast": lp + 8 * n_cand,                # read every level once, write candidates (xy + score)
        "quadtree": 12 * n_cand + 8 * n_kp,     # read candidates (xy, score, node), write survivors
        "blur": 2 * lp,                         # read + write every level
        "describe": (60 + 749 + 1369) * n_kp,   # per keypoint: 749-px disc + 37x37 blurred patch in, 60 B out
    }


def ncu_entry(kernel: str, frames: int):
    """the committed ncu --set full summary of `kernel` at this launch shape (profiles/r2_traffic.json, else
    profiles/r1_traffic.json; written by tools/ncu_traffic.py), or None"""
    for name in ("r2_traffic.json", "r1_traffic.json"):
        try:
            e = json.loads((ROOT / "profiles" / name).read_text())[kernel]
            if int(e["frames_per_launch"]) == frames:
                return dict(e, source=f"profiles/{name}")
        except Exception:
            continue
    return None


def ncu_traffic(kernel: str, frames: int):
    """dram bytes per launch of `kernel` from the committed ncu --set full capture, if it was taken at this launch shape"""
    e = ncu_entry(kernel, frames)
    return float(e["dram_bytes"]) if e and "dram_bytes" in e else None


def issue_roofline(kernel: str, frames: int, kernel_ms: float, sm_mhz: float, n_sm: int = 148) -> dict:
    """The path is instruction-issue bound, not HBM bound: warp instructions of one launch (ncu: sm__inst_executed.sum of
    the committed capture -- a property of the code and the input, not of the run) over the issue slots the launch had
    LIVE (SMs x 4 schedulers x measured SM clock x event-timed duration)."""
    e = ncu_entry(kernel, frames)
    if not e or "warp_instructions" not in e or not sm_mhz or not kernel_ms:
        return {"issue_frac": None}
    slots = n_sm * 4 * sm_mhz * 1e6 * kernel_ms / 1e3
    return {"issue_frac": float(e["warp_instructions"]) / slots, "warp_instructions_per_launch": float(e["warp_instructions"]),
            "issue_slots_busy_pct_under_ncu": e.get("issue_slots_busy_pct"), "issue_source": e["source"],
            "issue_how": "warp instructions per launch (ncu capture) / (148 SMs x 4 schedulers x SM clock sampled during the "
                         "timed region x this run's event-timed kernel duration)"}


def prestep_section(ORBextractor, synth, local, peak):
    """colour conversion and stereo rectification in front of the pyramid (K0).
    Device-resident raw frames; K0 is timed by events inside the real calls.  Algorithmic bytes per
    output pixel: C source bytes + 1 written (+ 6 map bytes when rectifying, L2-resident across
    frames: the map is shared by every frame of the batch)."""
    import torch
    out = {"reference": "src/Tracking.cc:176-201 (cvtColor), Examples/Stereo/stereo_euroc.cc:136-137 (remap)",
           "parity": "bit-exact against oracle/cvprim_pre.c, which is pinned to cv2 4.13.0"}
    n = 128

    def run(name, frames, fmt, maps, nfeat):
        dev = torch.from_numpy(frames).cuda(local)
        rows, cols = frames.shape[1:3]
        ch = 1 if frames.ndim == 3 else frames.shape[3]
        ext = ORBextractor(nfeat, SCALE, NLEVELS, INI_TH, MIN_TH, device=local, max_batch=n, want_pyramid=False)
        ext.set_input_format(fmt)
        if maps is not None:
            ext.set_rectify_map(*maps)
        ext.set_stage_timing(True)
        ms, tot = [], []
        for it in range(8):
            ext.timer_start()
            ext.extract_device(dev.data_ptr(), n, cols, rows)
            ext.timer_stop()
            ext.sync()
            if it >= 3:
                ms.append(ext.last_prestep_ms())
                tot.append(ext.timer_ms())
        k0 = float(np.median(ms))
        opix = (maps[0].size if maps is not None else rows * cols)
        alg = n * (rows * cols * ch + opix)
        out[name] = {"frames": n, "k0_ms": k0, "call_ms": float(np.median(tot)), "frames_per_s": n / (float(np.median(tot)) / 1e3),
                     "k0_GBps": alg / (k0 / 1e3) / 1e9, "k0_frac_of_hbm_peak": alg / (k0 / 1e3) / 1e9 / peak,
                     "algorithmic_bytes": alg}
        del ext, dev

    kitti = np.stack([synth.make_color_frame((ROWS, COLS), seed=7, frame=f % 8) for f in range(8)])
    run("bgr_to_gray_kitti", np.ascontiguousarray(kitti[np.arange(n) % 8]), "bgr", None, NFEATURES)
    euroc = np.stack([synth.make_frame(synth.EUROC_SHAPE, seed=40, frame=f) for f in range(8)])
    maps = synth.rectify_maps(*synth.EUROC[0], size=(752, 480))
    run("rectify_euroc_gray", np.ascontiguousarray(euroc[np.arange(n) % 8]), "gray", maps, 1200)
    return out


def poststep_section(ORBextractor, synth, local):
    """keypoint undistortion + 64x48 grid (K8) appended to the call.  K8 has no
    stage slot of its own: its time is the growth of the describe stage when a camera is set."""
    import torch
    K, dist = (517.306408, 516.469215, 318.643040, 255.313989), (0.262383, -0.953104, -0.005358, 0.002628, 1.163314)
    n = 128
    frames = np.stack([synth.make_frame(synth.TUM_SHAPE, seed=51, frame=f) for f in range(8)])
    frames = np.ascontiguousarray(frames[np.arange(n) % 8])
    dev = torch.from_numpy(frames).cuda(local)
    ext = ORBextractor(1000, SCALE, NLEVELS, INI_TH, MIN_TH, device=local, max_batch=n, want_pyramid=False)
    ext.set_stage_timing(True)
    res = {}
    for cam in (None, K):
        ext.set_camera(cam, dist if cam else ())
        d, tot = [], []
        for it in range(8):
            ext.timer_start()
            ext.extract_device(dev.data_ptr(), n, 640, 480)
            ext.timer_stop()
            ext.sync()
            if it >= 3:
                d.append(ext.last_stage_ms()["describe"])
                tot.append(ext.timer_ms())
        res["with" if cam else "without"] = (float(np.median(d)), float(np.median(tot)))
    out = {"reference": "src/Frame.cc:542-572, :574-609, :292-307", "config": "TUM1.yaml camera, 640x480, 1000 kp, 128 frames",
           "parity": "bit-exact against oracle/frame_oracle.c (undistortPoints pinned to cv2 4.13.0; whole step pinned to the reference's own Frame.cc, oracle/_ref/libreframe.so)",
           "k8_ms_per_128_frames": res["with"][0] - res["without"][0],
           "call_ms_without": res["without"][1], "call_ms_with": res["with"][1]}
    try:
        from oracle import pyoracle
        kps, _ = ext.fetch()[0]
        t1 = time.perf_counter()
        for _ in range(200):
            pyoracle.frame_post(kps, 640, 480, K, dist)
        out["cpu_us_per_frame_one_core"] = (time.perf_counter() - t1) / 200 * 1e6
        un, start, idx, b = ext.frame_grid(0)
        wun, wstart, widx, wb = pyoracle.frame_post(kps, 640, 480, K, dist)
        out["parity_check"] = "bit-exact" if (un.tobytes() == wun.tobytes() and np.array_equal(start, wstart)
                                              and np.array_equal(idx, widx)) else "MISMATCH"
    except Exception as e:
        out["parity_check"] = f"unavailable ({type(e).__name__})"
    return out


def tum_section(ORBextractor, synth, local):
    """BASELINE.json configs[0] / configs[1]: TUM shape (640x480), 1000 keypoints (TUM1.yaml) and the 2000-keypoint
    monocular-initialisation extractor -- device-resident throughput (512 frames per call, CUDA events) and the
    single-frame host-call latency, beside the headline KITTI numbers."""
    import torch
    F = 512
    uniq = np.stack([synth.make_frame(synth.TUM_SHAPE, seed=7, frame=f) for f in range(UNIQUE)])
    batch = np.ascontiguousarray(uniq[np.arange(F) % UNIQUE])
    dev = torch.from_numpy(batch).cuda(local)
    out = {}
    for name, nf in (("T1k", 1000), ("T2k", 2000)):
        ext = ORBextractor(nf, SCALE, NLEVELS, INI_TH, MIN_TH, device=local, max_batch=F, want_pyramid=False)
        for _ in range(3):
            ext.extract_device(dev.data_ptr(), F, 640, 480)
        ext.sync()
        ext.timer_start()
        for _ in range(5):
            ext.extract_device(dev.data_ptr(), F, 640, 480)
        ext.timer_stop()
        ext.sync()
        ms = ext.timer_ms() / 5
        res = ext.fetch()
        one = ORBextractor(nf, SCALE, NLEVELS, INI_TH, MIN_TH, device=local, max_batch=1, want_pyramid=False)
        img = batch[0].copy()
        ts = []
        for it in range(220):
            t1 = time.perf_counter()
            one._lib.borb_extract_host(one._h, img.ctypes.data, 1, 640, 480, 640, 0, 0)
            if it >= 20:
                ts.append((time.perf_counter() - t1) * 1e6)
        e = {"nfeatures": nf, "frames_per_s_device_resident": F / (ms / 1e3), "ms_per_512_frames": ms,
             "keypoints_per_frame": float(np.mean([len(k) for k, _ in res[:UNIQUE]])), "latency_p50_us": float(np.percentile(ts, 50))}
        try:
            from oracle.pyoracle import Oracle
            ok, od = Oracle(nf, SCALE, NLEVELS, INI_TH, MIN_TH).extract(batch[0])
            e["parity_check"] = "bit-exact" if (res[0][0].tobytes() == ok.tobytes() and res[0][1].tobytes() == od.tobytes()) else "MISMATCH"
        except Exception as ex:
            e["parity_check"] = f"unavailable ({type(ex).__name__})"
        out[name] = e
    return out


def matcher_section(ORBextractor, synth, local):
    """SURVEY.md 8(f)-4, first consumer: the tracking thread's frame-to-frame matcher
    ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) on the device
    (borb_search_by_projection).  Scenario: two consecutive KITTI-shape stereo frames extracted and
    stereo-matched on the GPU; the last frame's stereo depths become map points (tests/matchcase.py);
    the current frame's keypoints, descriptors, grid and mvuRight stay on the device.  The timed call is
    the whole host call (upload of the last frame's 53 B per keypoint, two kernels, download of the matches)."""
    sys.path.insert(0, str(ROOT / "tests"))
    import matchcase
    K, BF = matchcase.KITTI_K, matchcase.KITTI_BF
    mb = BF / K[0]

    def gpu_frame(frame):
        l = ORBextractor(NFEATURES, SCALE, NLEVELS, INI_TH, MIN_TH, device=local)
        r = ORBextractor(NFEATURES, SCALE, NLEVELS, INI_TH, MIN_TH, device=local)
        l.set_camera(K, (0.0, 0.0, 0.0, 0.0))
        l(synth.make_frame((ROWS, COLS), seed=71, frame=frame, eye=0))
        r(synth.make_frame((ROWS, COLS), seed=71, frame=frame, eye=1))
        l.stereo_match(r, BF, mb)
        kps, desc = l.result(0)
        u, z, _ = l.stereo_result(0)
        un, start, idx, bounds = l.frame_grid(0)
        return l, r, {"kps": kps, "desc": desc, "un_xy": un, "u_right": u, "depth": z, "cell_start": start,
                      "cell_idx": idx, "bounds": bounds}
    _, _, last = gpu_frame(0)
    ext, right, cur = gpu_frame(1)
    T_last = matchcase.pose(0.01, -0.02, 0.005, (0.3, -0.1, 0.2))
    cur["Tcw"] = matchcase.compose(matchcase.pose(rx=1.0 / K[1], ry=-3.0 / K[0]), T_last)
    last = matchcase.map_points(last, T_last, K, np.random.default_rng(71))
    lf = dict(last)
    lf["octave"], lf["angle"] = last["kps"]["octave"], last["kps"]["angle"]
    out = {"reference": "src/ORBmatcher.cc:1503-1668 (+ Frame::GetFeaturesInArea src/Frame.cc:421-508)",
           "parity": "identical mvpMapPoints and nmatches against oracle/match_oracle.cpp, which is pinned to the reference's own ORBmatcher.cc (oracle/_ref/librefmatch.so)",
           "config": "KITTI stereo frame pair, 2000 kp, %d map points in the last frame" % int((last["mp_flags"] & 1).sum())}
    scale = ext.GetScaleFactors()
    cam = np.array([K[0], K[1], K[2], K[3], BF, mb], np.float32)
    for th in (7.0, 15.0):
        for _ in range(10):
            m, nm = ext.search_by_projection(lf, cur["Tcw"], BF, mb, th, use_stereo=True)
        ts = []
        for _ in range(100):
            t1 = time.perf_counter()
            m, nm = ext.search_by_projection(lf, cur["Tcw"], BF, mb, th, use_stereo=True)
            ts.append((time.perf_counter() - t1) * 1e6)
        e = {"call_p50_us": float(np.percentile(ts, 50)), "nmatches": nm, "rounds": ext.match_rounds(),
             "rescans": ext.match_rescans()}
        try:
            from oracle import pyoracle
            tc = []
            for _ in range(20):
                t1 = time.perf_counter()
                w, wnm = pyoracle.search_by_projection(cur, last, cam, cur["bounds"], scale, th)
                tc.append((time.perf_counter() - t1) * 1e6)
            e["cpu_p50_us_one_core"] = float(np.percentile(tc, 50))
            e["parity_check"] = "identical" if (np.array_equal(m, w) and nm == wnm) else "MISMATCH"
        except Exception as ex:
            e["parity_check"] = f"unavailable ({type(ex).__name__})"
        out["th_%d" % int(th)] = e
    # the local-map matcher of Tracking::SearchLocalPoints on the same current frame: the motion-model matches
    # occupy their slots, the local map = the unmatched map points plus two jittered copies (tests/matchcase.py)
    try:
        from oracle import pyoracle
        m7, _ = ext.search_by_projection(lf, cur["Tcw"], BF, mb, 7.0, use_stereo=True)
        occ = np.zeros(len(m7), np.uint8)
        held = m7 >= 0
        occ[held] = np.where(last["mp_flags"][m7[held]] & 2, 2, 1)
        rng = np.random.default_rng(72)
        copies = 3
        matched = np.zeros(len(last["kps"]), bool)
        matched[m7[held]] = True
        xyz = np.concatenate([last["mp_xyz"]] + [last["mp_xyz"] + rng.normal(0, 0.03, last["mp_xyz"].shape).astype(np.float32)
                                                 for _ in range(copies - 1)]).astype(np.float64)
        has = np.concatenate([(last["mp_flags"] & 1).astype(bool) & ~matched] + [(last["mp_flags"] & 1).astype(bool)] * (copies - 1))
        T = cur["Tcw"].astype(np.float64)
        Xc = xyz @ T[:, :3].T + T[:, 3]
        z = np.where(np.abs(Xc[:, 2]) > 1e-6, Xc[:, 2], 1.0)
        u, v = K[0] * Xc[:, 0] / z + K[2], K[1] * Xc[:, 1] / z + K[3]
        b = cur["bounds"]
        inside = (Xc[:, 2] > 0) & (u > b[0]) & (u < b[1]) & (v > b[2]) & (v < b[3])
        N = len(xyz)
        proj = np.stack([u, v, u - BF / z, rng.uniform(0.99, 1.0, N)], 1).astype(np.float32)
        mps = {"proj": proj, "desc": np.concatenate([last["mp_desc"]] * copies),
               "flags": ((has & inside).astype(np.uint8) | ((rng.random(N) < 0.9).astype(np.uint8) << 1)),
               "level": np.clip(np.concatenate([last["kps"]["octave"]] * copies) + rng.integers(-1, 2, N), 0, 7).astype(np.int32)}
        loc = {"reference": "src/ORBmatcher.cc:59-154 (Tracking::SearchLocalPoints)",
               "config": "%d local map points, %d in view, %d slots held" % (N, int((mps["flags"] & 1).sum()), int((occ == 2).sum()))}
        for th in (1.0, 3.0):
            for _ in range(10):
                m, nm = ext.search_local_points(mps, occ, th, 0.8, use_stereo=True)
            ts = []
            for _ in range(100):
                t1 = time.perf_counter()
                m, nm = ext.search_local_points(mps, occ, th, 0.8, use_stereo=True)
                ts.append((time.perf_counter() - t1) * 1e6)
            tc = []
            for _ in range(20):
                t1 = time.perf_counter()
                w, wnm = pyoracle.search_local_points(cur, cur["bounds"], scale, mps, occ, th, 0.8)
                tc.append((time.perf_counter() - t1) * 1e6)
            loc["th_%d" % int(th)] = {"call_p50_us": float(np.percentile(ts, 50)), "nmatches": nm, "rounds": ext.match_rounds(),
                                      "rescans": ext.match_rescans(), "cpu_p50_us_one_core": float(np.percentile(tc, 50)),
                                      "parity_check": "identical" if (np.array_equal(m, w) and nm == wnm) else "MISMATCH"}
        out["local_map"] = loc
    except Exception as ex:
        out["local_map"] = f"unavailable ({type(ex).__name__}: {ex})"
    # the relocalisation matcher ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist): the last
    # frame as the keyframe, its map points with the scale-invariance distances of their own observation, 20 % of them
    # already found and 15 % of the current slots set (tests/matchcase.py: make_reloc_case builds the same on the CPU)
    try:
        from oracle import pyoracle
        import ctypes
        import ctypes.util
        libm = ctypes.CDLL(ctypes.util.find_library("m") or "libm.so.6")
        libm.logf.restype, libm.logf.argtypes = ctypes.c_float, [ctypes.c_float]
        rng = np.random.default_rng(73)
        n = len(last["mp_flags"])
        Tl = T_last.astype(np.float64)
        Ow = -Tl[:, :3].T @ Tl[:, 3]
        d = np.linalg.norm(last["mp_xyz"].astype(np.float64) - Ow, axis=1)
        sc = np.asarray(scale, np.float64)
        max_d = d * sc[last["kps"]["octave"]] * rng.uniform(0.75, 1.25, n)
        kf = {"flags": ((last["mp_flags"] & 1) | ((rng.random(n) < 0.05).astype(np.uint8) << 1)
                        | ((rng.random(n) < 0.2).astype(np.uint8) << 2)).astype(np.uint8),
              "xyz": last["mp_xyz"], "desc": last["mp_desc"], "dist": np.stack([max_d / sc[-1], max_d], 1).astype(np.float32),
              "angle": last["kps"]["angle"].astype(np.float32)}
        occ_r = (rng.random(len(cur["kps"])) < 0.15).astype(np.uint8)
        rel = {"reference": "src/ORBmatcher.cc:1670-1798 (Tracking::Relocalization), MapPoint::PredictScale src/MapPoint.cc:575-590",
               "config": "%d keyframe map points, %d already found, %d current slots set" % (
                   int((kf["flags"] & 1).sum()), int(((kf["flags"] & 5) == 5).sum()), int(occ_r.sum()))}
        for th, od in ((10.0, 100), (3.0, 64)):
            for _ in range(10):
                m, nm = ext.search_by_projection_keyframe(kf, cur["Tcw"], occ_r, th, od)
            ts = []
            for _ in range(100):
                t1 = time.perf_counter()
                m, nm = ext.search_by_projection_keyframe(kf, cur["Tcw"], occ_r, th, od)
                ts.append((time.perf_counter() - t1) * 1e6)
            tc = []
            for _ in range(20):
                t1 = time.perf_counter()
                w, wnm = pyoracle.search_by_projection_kf(cur, cam, cur["bounds"], scale, libm.logf(float(np.float32(SCALE))), kf,
                                                          occ_r, th, od, True)
                tc.append((time.perf_counter() - t1) * 1e6)
            rel["th_%d_dist_%d" % (int(th), od)] = {
                "call_p50_us": float(np.percentile(ts, 50)), "nmatches": nm, "rounds": ext.match_rounds(),
                "rescans": ext.match_rescans(), "cpu_p50_us_one_core": float(np.percentile(tc, 50)),
                "parity_check": "identical" if (np.array_equal(m, w) and nm == wnm) else "MISMATCH"}
        out["relocalisation"] = rel
    except Exception as ex:
        out["relocalisation"] = f"unavailable ({type(ex).__name__}: {ex})"
    # the matchers of the back-end threads that project map points into a KEYFRAME (loaded into a handle with
    # borb_load_frame): Fuse(pKF, vpMapPoints, th) of LocalMapping::SearchInNeighbors and SearchBySim3 of LoopClosing
    try:
        from oracle import pyoracle

        def timed(fn, reps):
            for _ in range(max(2, reps // 10)):
                r = fn()
            ts = []
            for _ in range(reps):
                t1 = time.perf_counter()
                r = fn()
                ts.append((time.perf_counter() - t1) * 1e6)
            return r, float(np.percentile(ts, 50))
        kf_, cam_, bounds_, scale_, inv_s2, log_s, mps_ = matchcase.make_fuse_case(seed=74)
        hk = ORBextractor(NFEATURES, SCALE, NLEVELS, INI_TH, MIN_TH, device=local)
        hk.set_camera(K, (0.0, 0.0, 0.0, 0.0))
        _, t_load = timed(lambda: hk.load_frame(kf_["kps"], kf_["desc"], COLS, ROWS, kf_["u_right"]), 50)
        Ow_ = pyoracle.camera_centre(kf_["Tcw"])
        (b_, nf_), t_gpu = timed(lambda: hk.fuse_search(mps_, kf_["Tcw"], Ow_, float(cam_[4]), 3.0, use_stereo=True), 100)
        (w_, wnf_), t_cpu = timed(lambda: pyoracle.fuse(kf_, cam_, bounds_, scale_, inv_s2, log_s, mps_, 3.0), 10)
        out["fuse"] = {"reference": "src/ORBmatcher.cc:968-1131 (LocalMapping::SearchInNeighbors): the search; the map update stays with the caller",
                       "config": "%d map points into a stereo keyframe of %d keypoints" % (int((mps_["flags"] & 1).sum()), len(kf_["kps"])),
                       "load_keyframe_p50_us": t_load, "call_p50_us": t_gpu, "nfused": int(nf_), "cpu_p50_us_one_core": t_cpu,
                       "parity_check": "identical" if (np.array_equal(b_, w_) and nf_ == wnf_) else "MISMATCH"}
        k1_, k2_, cam_, bounds_, scale_, inv_s2, log_s, p1_, p2_, pre_, s12_, R12_, t12_ = matchcase.make_sim3_pair_case(seed=75)
        h1_ = ORBextractor(NFEATURES, SCALE, NLEVELS, INI_TH, MIN_TH, device=local)
        h2_ = ORBextractor(NFEATURES, SCALE, NLEVELS, INI_TH, MIN_TH, device=local)
        for h_, k_ in ((h1_, k1_), (h2_, k2_)):
            h_.set_camera(K, (0.0, 0.0, 0.0, 0.0))
            h_.load_frame(k_["kps"], k_["desc"], COLS, ROWS)
        T12_, T21_ = pyoracle.sim3_relative(s12_, R12_, t12_)
        (m_, nf_), t_gpu = timed(lambda: h1_.search_by_sim3(h2_, p1_, p2_, k1_["Tcw"], k2_["Tcw"], T12_, T21_, 7.5, pre_), 100)
        (w_, wnf_), t_cpu = timed(lambda: pyoracle.search_by_sim3(k1_, k2_, cam_, bounds_, scale_, log_s, p1_, p2_, pre_, s12_, R12_, t12_, 7.5), 10)
        out["search_by_sim3"] = {"reference": "src/ORBmatcher.cc:1262-1501 (LoopClosing::ComputeSim3)", "call_p50_us": t_gpu,
                                 "nfound": int(nf_), "cpu_p50_us_one_core": t_cpu,
                                 "parity_check": "identical" if (np.array_equal(m_, w_) and nf_ == wnf_) else "MISMATCH"}
        for h_ in (hk, h1_, h2_):
            h_.close()
    except Exception as ex:
        out["keyframe_matchers"] = f"unavailable ({type(ex).__name__}: {ex})"
    # the whole per-frame front end of the tracking thread for a stereo frame, as one sequence through the host ABI:
    # both eyes extracted on two threads (+ undistortion / grid), stereo association, frame-to-frame matcher,
    # local-map matcher -- the current frame's keypoints, descriptors, grid and mvuRight never leave the device
    # in between (the 28+32 B per keypoint results are still downloaded by the extract calls, as the shim needs them)
    try:
        import threading
        imgs = [synth.make_frame((ROWS, COLS), seed=71, frame=1, eye=e) for e in (0, 1)]
        pair = [ext, right]

        def eye(i):
            pair[i]._lib.borb_extract_host(pair[i]._h, imgs[i].ctypes.data, 1, COLS, ROWS, COLS, 0, 0)
            pair[i]._n = 1
        mps_ok = isinstance(out.get("local_map"), dict)
        ts = []
        for it in range(120):
            t1 = time.perf_counter()
            th_ = threading.Thread(target=eye, args=(1,))
            th_.start()
            eye(0)
            th_.join()
            ext.stereo_match(right, BF, mb, keep_on_device=True)
            m7, _ = ext.search_by_projection(lf, cur["Tcw"], BF, mb, 7.0, use_stereo=True)
            if mps_ok:
                ext.search_local_points(mps, occ, 1.0, 0.8, use_stereo=True)
            if it >= 20:
                ts.append((time.perf_counter() - t1) * 1e6)
        out["stereo_front_end_p50_us"] = float(np.percentile(ts, 50))
        out["stereo_front_end"] = "extract L+R (2 threads) + undistort/grid + stereo association + SearchByProjection(th 7) + SearchLocalPoints(th 1), one KITTI stereo frame, host call sequence"
    except Exception as ex:
        out["stereo_front_end_p50_us"] = f"unavailable ({type(ex).__name__}: {ex})"
    # the monocular-initialisation matcher (ORBmatcher::SearchForInitialization, window 100) on two TUM-shape frames of
    # the 2 x nFeatures initialisation extractor
    try:
        from oracle import pyoracle
        Kt, Dt = matchcase.TUM1_K, matchcase.TUM1_D
        ini = ORBextractor(2000, SCALE, NLEVELS, INI_TH, MIN_TH, device=local)
        ini.set_camera(Kt, Dt)
        frames = {}
        for f in (0, 1):
            kps, desc = ini(synth.make_frame(synth.TUM_SHAPE, seed=73, frame=f))
            un, start, idx, b = ini.frame_grid(0)
            frames[f] = {"kps": kps, "desc": desc, "un_xy": un, "cell_start": start, "cell_idx": idx, "bounds": b}
        f1, f2 = frames[0], frames[1]
        prev = f1["un_xy"].copy()
        for _ in range(10):
            m, pv, nm = ini.search_for_initialization(f1, prev, 100, 0.9)
        ts = []
        for _ in range(100):
            t1 = time.perf_counter()
            m, pv, nm = ini.search_for_initialization(f1, prev, 100, 0.9)
            ts.append((time.perf_counter() - t1) * 1e6)
        tc = []
        for _ in range(10):
            t1 = time.perf_counter()
            w, wpv, wnm = pyoracle.search_for_initialization(f1, f2, f2["bounds"], prev, 100, 0.9, True)
            tc.append((time.perf_counter() - t1) * 1e6)
        out["initialisation"] = {"reference": "src/ORBmatcher.cc:487-632 (Tracking::MonocularInitialization)",
                                 "config": "TUM 640x480, 2000 kp per frame, %d level-0 queries, window 100" % int((f1["kps"]["octave"] == 0).sum()),
                                 "call_p50_us": float(np.percentile(ts, 50)), "nmatches": nm, "rounds": ini.match_rounds(),
                                 "rescans": ini.match_rescans(), "cpu_p50_us_one_core": float(np.percentile(tc, 50)),
                                 "parity_check": "identical" if (np.array_equal(m, w) and nm == wnm and np.array_equal(pv, wpv)) else "MISMATCH"}
    except Exception as ex:
        out["initialisation"] = f"unavailable ({type(ex).__name__}: {ex})"
    return out


def bow_section(ORBextractor, synth, local):
    """Frame::ComputeBoW (src/Frame.cc:532-539 -> DBoW2 transform, levelsup 4) and
    ORBmatcher::SearchByBoW(KeyFrame*, Frame&) (src/ORBmatcher.cc:195-351) on the device.  The reference's vocabulary
    blob (ORBvoc.txt, k = 10, L = 6) is not in the reference tree: a SYNTHETIC vocabulary of the same shape
    (tests/golden/make_vocabulary.py, k = 10, L = 5: 111 110 nodes -- L = 6 is 145 MB of text) in the reference's text
    format.  Timed: the whole host calls, on a KITTI-shape frame whose descriptors are already on the device."""
    import tempfile
    sys.path.insert(0, str(ROOT / "tests" / "golden"))
    import make_vocabulary
    from orb_slam2_test_b200 import ORBVocabulary
    out = {"reference": "src/Frame.cc:532-539, Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1270, src/ORBmatcher.cc:195-351",
           "vocabulary": "synthetic, k = 10, L = 5, L1 scoring, TF-IDF weighting (the real ORBvoc.txt is absent from the reference tree)"}
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "voc.txt")
        make_vocabulary.write_vocabulary(path, 10, 5, seed=5)
        voc = ORBVocabulary(path, device=local)
        ext = ORBextractor(NFEATURES, SCALE, NLEVELS, INI_TH, MIN_TH, device=local)
        kk, kd = ext(synth.make_frame((ROWS, COLS), seed=81, frame=0))
        kk, kd = kk.copy(), kd.copy()
        fk, fd = ext(synth.make_frame((ROWS, COLS), seed=81, frame=1))
        for _ in range(10):
            bow = ext.compute_bow(voc)
        ts = []
        for _ in range(100):
            t1 = time.perf_counter()
            bow = ext.compute_bow(voc)
            ts.append((time.perf_counter() - t1) * 1e6)
        out["compute_bow"] = {"call_p50_us": float(np.percentile(ts, 50)), "features": int(len(fd)), "words": int(len(bow["word_id"])),
                              "nodes": int(len(bow["node_id"]))}
        rng = np.random.default_rng(81)
        mp = (rng.random(len(kk)) < 0.7).astype(np.uint8)
        kf = {"desc": kd, "angle": kk["angle"], "mp_flags": mp, "fv": voc.transform(kd, 4)}
        for _ in range(10):
            m, nm = ext.search_by_bow(kf, 0.7)
        ts = []
        for _ in range(100):
            t1 = time.perf_counter()
            m, nm = ext.search_by_bow(kf, 0.7)
            ts.append((time.perf_counter() - t1) * 1e6)
        out["search_by_bow"] = {"call_p50_us": float(np.percentile(ts, 50)), "nmatches": int(nm), "keyframe_map_points": int(mp.sum()),
                                "nodes": int(len(kf["fv"]["node_id"])),
                                "note": "levelsup 4 on this L = 5 tree groups the features under 10 nodes (up to 480 per node); the walk "
                                        "inside a node is sequential, as in the reference"}
        # the granularity of the real vocabulary (L = 6, levelsup 4 -> 100 nodes): levelsup 3 on this tree
        kf3 = dict(kf, fv=voc.transform(kd, 3))
        ext.compute_bow(voc, levelsup=3)
        for _ in range(10):
            m3, nm3 = ext.search_by_bow(kf3, 0.7)
        ts = []
        for _ in range(100):
            t1 = time.perf_counter()
            m3, nm3 = ext.search_by_bow(kf3, 0.7)
            ts.append((time.perf_counter() - t1) * 1e6)
        out["search_by_bow_100_nodes"] = {"call_p50_us": float(np.percentile(ts, 50)), "nmatches": int(nm3),
                                          "nodes": int(len(kf3["fv"]["node_id"])), "levelsup": 3}
        bow = ext.compute_bow(voc)
        try:
            from oracle import pyoracle
            want = pyoracle.BowVocabulary(path)
            tc = []
            for _ in range(10):
                t1 = time.perf_counter()
                wb = want.transform(fd, 4)
                tc.append((time.perf_counter() - t1) * 1e6)
            out["compute_bow"]["cpu_p50_us_one_core"] = float(np.percentile(tc, 50))
            same = all(np.array_equal(np.asarray(bow[k]).view(np.uint8), np.asarray(wb[k]).view(np.uint8)) for k in wb)
            out["compute_bow"]["parity_check"] = "bit-exact" if same else "MISMATCH"
            kfv = want.transform(kd, 4)
            tc = []
            for _ in range(10):
                t1 = time.perf_counter()
                w, wnm = want.search_by_bow(kk, kd, mp, fk, fd, 0.7, True, kf_fv=kfv, f_fv=wb)
                tc.append((time.perf_counter() - t1) * 1e6)
            out["search_by_bow"]["cpu_p50_us_one_core"] = float(np.percentile(tc, 50))
            out["search_by_bow"]["parity_check"] = "identical" if (np.array_equal(m, w) and nm == wnm) else "MISMATCH"
            wb3, kfv3 = want.transform(fd, 3), want.transform(kd, 3)
            tc = []
            for _ in range(10):
                t1 = time.perf_counter()
                w3, wnm3 = want.search_by_bow(kk, kd, mp, fk, fd, 0.7, True, kf_fv=kfv3, f_fv=wb3)
                tc.append((time.perf_counter() - t1) * 1e6)
            out["search_by_bow_100_nodes"]["cpu_p50_us_one_core"] = float(np.percentile(tc, 50))
            out["search_by_bow_100_nodes"]["parity_check"] = "identical" if (np.array_equal(m3, w3) and nm3 == wnm3) else "MISMATCH"
        except Exception as ex:
            out["parity_check"] = f"unavailable ({type(ex).__name__}: {ex})"
        # the back-end threads' bag-of-words matchers between two keyframes (both sides uploaded): SearchByBoW(pKF1, pKF2)
        # of LoopClosing::ComputeSim3 and SearchForTriangulation of LocalMapping::CreateNewMapPoints
        try:
            from oracle import pyoracle
            sys.path.insert(0, str(ROOT / "tests"))
            import bowcase
            want = pyoracle.BowVocabulary(path)
            mp2 = (rng.random(len(fk)) < 0.7).astype(np.uint8)
            kf1 = {"desc": kd, "angle": kk["angle"], "mp_flags": mp, "fv": voc.transform(kd, 3)}
            kf2 = {"desc": fd, "angle": fk["angle"], "mp_flags": mp2, "fv": voc.transform(fd, 3)}
            fv1, fv2 = want.transform(kd, 3), want.transform(fd, 3)

            def timed(fn, reps):
                for _ in range(max(2, reps // 10)):
                    r = fn()
                ts = []
                for _ in range(reps):
                    t1 = time.perf_counter()
                    r = fn()
                    ts.append((time.perf_counter() - t1) * 1e6)
                return r, float(np.percentile(ts, 50))
            (m, nm), t_gpu = timed(lambda: voc.search_by_bow(kf1, kf2, 0.75), 100)
            (w, wnm), t_cpu = timed(lambda: want.search_by_bow_kf(kk, kd, mp, fk, fd, mp2, 0.75, True, fv1=fv1, fv2=fv2), 10)
            out["search_by_bow_keyframes"] = {"reference": "src/ORBmatcher.cc:634-777", "call_p50_us": t_gpu, "nmatches": int(nm),
                                              "cpu_p50_us_one_core": t_cpu, "nodes": int(len(fv1["node_id"])),
                                              "parity_check": "identical" if (np.array_equal(m, w) and nm == wnm) else "MISMATCH"}
            a, b, geo = bowcase.triangulation_case(shape=(ROWS, COLS), seed=81, t=0, nf=NFEATURES, baseline=(0.3, 0.0, 0.05),
                                                   K=(718.856, 718.856, 607.1928, 185.2157), has=0.6)
            t1_ = {"desc": a["desc"], "angle": a["kps"]["angle"], "mp_flags": a["mp"], "fv": voc.transform(a["desc"], 3),
                   "xy": np.stack([a["kps"]["x"], a["kps"]["y"]], 1).astype(np.float32), "octave": a["kps"]["octave"].astype(np.int32)}
            t2_ = {"desc": b["desc"], "angle": b["kps"]["angle"], "mp_flags": b["mp"], "fv": voc.transform(b["desc"], 3),
                   "xy": np.stack([b["kps"]["x"], b["kps"]["y"]], 1).astype(np.float32), "octave": b["kps"]["octave"].astype(np.int32)}
            g = dict(geo, Ow1=pyoracle.camera_centre(a["Tcw"]), Tcw2=b["Tcw"])
            f1_, f2_ = want.transform(a["desc"], 3), want.transform(b["desc"], 3)
            (m, nm), t_gpu = timed(lambda: voc.search_for_triangulation(t1_, t2_, g), 100)
            (w, wnm), t_cpu = timed(lambda: want.search_for_triangulation(a, b, geo, False, True, fv1=f1_, fv2=f2_), 10)
            out["search_for_triangulation"] = {"reference": "src/ORBmatcher.cc:779-966 (+ CheckDistEpipolarLine :165-182)",
                                               "call_p50_us": t_gpu, "nmatches": int(nm), "cpu_p50_us_one_core": t_cpu,
                                               "parity_check": "identical" if (np.array_equal(m, w) and nm == wnm) else "MISMATCH"}
        except Exception as ex:
            out["keyframe_matchers"] = f"unavailable ({type(ex).__name__}: {ex})"
        ext.close()
        voc.close()
    return out


def stereo_section(ORBextractor, synth, local, two, dev):
    """Frame::ComputeStereoMatches on the device (borb_stereo_match).  Reported
    beside the headline, not part of `value`: (1) a stereo pair through the host ABI with the match on
    the device and NO pyramid download, against the same pair with both pyramids brought back for a
    host-side match; (2) batched pairs, device resident: both eyes' extraction + match per pair;
    (3) the CPU restatement of the match alone on one host core."""
    import threading
    BF, FX = 386.1448, 718.856
    mb = BF / FX
    out = {"reference": "src/Frame.cc:619-834", "parity": "bit-exact against oracle/stereo_oracle.cpp, which is pinned to the reference's own Frame.cc (oracle/_ref/libreframe.so)"}
    li = synth.make_frame((ROWS, COLS), seed=31, frame=0, eye=0)
    ri = synth.make_frame((ROWS, COLS), seed=31, frame=0, eye=1)
    imgs = [li, ri]

    def eye(i):
        two[i]._lib.borb_extract_host(two[i]._h, imgs[i].ctypes.data, 1, COLS, ROWS, COLS, 0, 0)
        two[i]._n = 1

    ts = []
    for it in range(220):
        t1 = time.perf_counter()
        th = threading.Thread(target=eye, args=(1,))
        th.start()
        eye(0)
        th.join()
        two[0].stereo_match(two[1], BF, mb)
        if it >= 20:
            ts.append((time.perf_counter() - t1) * 1e6)
    out["pair_extract_plus_device_match_p50_us"] = float(np.percentile(ts, 50))
    u, z, kept = two[0].stereo_result(0)
    out["matched_per_pair"] = kept
    # the match alone, CUDA events on the left handle's stream
    ms = []
    for _ in range(50):
        two[0].timer_start()
        two[0].stereo_match(two[1], BF, mb, keep_on_device=True)
        two[0].timer_stop()
        two[0].sync()
        ms.append(two[0].timer_ms())
    out["match_kernels_p50_us_one_pair"] = float(np.percentile(ms, 50) * 1e3)
    # CPU restatement of the match alone (inputs: this run's keypoints, descriptors, pyramids)
    try:
        from oracle import pyoracle
        for e in two:
            e.fetch(want_pyramid=True)
        kl, dl = two[0].result(0)
        kr, dr = two[1].result(0)
        pl = [two[0].pyramid_level(0, l) for l in range(NLEVELS)]
        pr = [two[1].pyramid_level(0, l) for l in range(NLEVELS)]
        t1 = time.perf_counter()
        for _ in range(20):
            wu, wz, _, wk = pyoracle.stereo_match(kl, dl, kr, dr, pl, pr, two[0].GetScaleFactors(),
                                                  two[0].GetInverseScaleFactors(), BF, mb)
        out["cpu_match_us_one_pair_one_core"] = (time.perf_counter() - t1) / 20 * 1e6
        out["parity_check"] = "bit-exact" if (wu.tobytes() == u.tobytes() and wz.tobytes() == z.tobytes()) else "MISMATCH"
    except Exception as e:
        out["parity_check"] = f"unavailable ({type(e).__name__})"
    # batched pairs, device resident: frames 2i / 2i+1 of the resident batch stand for the two eyes
    # (timing only -- these are not rectified pairs, so few of them match)
    P = max(1, min(128, dev.shape[0] // 2))
    lb = ORBextractor(NFEATURES, SCALE, NLEVELS, INI_TH, MIN_TH, device=local, max_batch=P, want_pyramid=False)
    rb = ORBextractor(NFEATURES, SCALE, NLEVELS, INI_TH, MIN_TH, device=local, max_batch=P, want_pyramid=False)
    lp = dev.data_ptr()
    rp = dev.data_ptr() + P * ROWS * COLS
    for rep in range(5):
        if rep == 2:
            lb.sync(); rb.sync()
            lb.timer_start()
        lb.extract_device(lp, P, COLS, ROWS)
        rb.extract_device(rp, P, COLS, ROWS)
        lb.stereo_match(rb, BF, mb, keep_on_device=True)
    lb.timer_stop()
    rb.sync(); lb.sync()
    out["batched_pairs_per_s_device_resident"] = 3 * P / (lb.timer_ms() / 1e3)
    ms = []
    for _ in range(5):
        lb.timer_start()
        lb.stereo_match(rb, BF, mb, keep_on_device=True)
        lb.timer_stop()
        lb.sync()
        ms.append(lb.timer_ms())
    out["match_ms_per_batch"] = float(np.median(ms))
    out["pairs_per_batch"] = P
    return out


def kseq_section(args, rank, world, local, synth, OrbStream, pinned_empty):
    """BASELINE.json configs[4] as stated: the 4541-pair synthetic KITTI-00-shaped stereo sequence, host-fed, pair f on
    rank f mod N, through one library feeder ring per GPU: left + right extraction, Frame::ComputeStereoMatches,
    UndistortKeyPoints + AssignFeaturesToGrid on the device; mvKeys, mDescriptors (both eyes), mvuRight, mvDepth, mvKeysUn
    and mGrid come back for every pair.  Timed pass: wall clock between barriers, max over ranks, generation excluded.
    Second (untimed) pass: per-pair SHA-256 digests, re-ordered by frame index on rank 0, digest of digests compared with
    tests/golden/kseq_digests.json = the reference's own Frame.cc over the same sequence (make_golden_kseq.py)."""
    import torch
    from orb_slam2_test_b200 import sequence
    n_pairs = int(os.environ.get("BENCH_KSEQ_PAIRS", str(synth.KSEQ_PAIRS)))
    if n_pairs <= 0:
        return None
    B = int(os.environ.get("BENCH_KSEQ_BATCH", "64"))
    mine = list(sequence.shard_indices(n_pairs, rank, world))
    t0 = time.perf_counter()
    left, right = pinned_empty((len(mine), ROWS, COLS)), pinned_empty((len(mine), ROWS, COLS))
    synth.kseq_fill(mine, left, right, threads=max(2, len(os.sched_getaffinity(0)) // world))
    gen_s = time.perf_counter() - t0
    K, BF = synth.KSEQ_K, synth.KSEQ_BF

    def stream():
        st = OrbStream(NFEATURES, SCALE, NLEVELS, INI_TH, MIN_TH, COLS, ROWS, batch=B, slots=3, device=local, stereo=True)
        st.set_stereo(BF, BF / K[0])
        st.set_camera(K, (0.0, 0.0, 0.0, 0.0))
        return st

    st = stream()
    warm = min(len(mine), 4 * B)
    sequence.run_stream_shard(st, mine[:warm], left, right, digest=False)  # graphs captured, buffers touched
    barrier(world)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    timed = sequence.run_stream_shard(st, mine, left, right, digest=False)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3
    barrier(world)
    ms_all = all_max(ms, world)
    up_b, down_b = st.link_bytes()
    launches = st.launches_per_push() * ((len(mine) + B - 1) // B)
    # second pass, untimed: digests
    res = sequence.run_stream_shard(st, mine, left, right, digest=True)
    st.close()
    del left, right
    merged = sequence.gather_items_to_rank0(res, rank, world)
    if rank != 0:
        return None
    out = {"workload": "Kseq: %d synthetic KITTI-00-shaped stereo pairs (1241x376, 2000 kp per eye), pair f -> rank f mod N" % n_pairs,
           "reference": "Examples/Stereo/stereo_kitti.cc:66-108 -> Frame::Frame src/Frame.cc:79-151",
           "pairs": n_pairs, "pairs_per_s": n_pairs / (ms_all / 1e3), "images_per_s": 2 * n_pairs / (ms_all / 1e3),
           "wall_ms": ms_all, "n_gpus": world, "batch_pairs": B,
           "h2d_bytes_per_pair": up_b // B, "d2h_bytes_per_pair": down_b // B,
           "link_GBps": (up_b + down_b) / B * n_pairs / (ms_all / 1e3) / 1e9,
           "gpu_launches_rank0": launches, "generation_s_excluded": gen_s,
           "keypoints_left_mean": float(np.mean([r.n_left for r in merged])),
           "matched_mean": float(np.mean([r.n_matched for r in merged])),
           "timed_pass_counts_equal_digest_pass": [(r.n_left, r.n_right, r.n_matched) for r in timed] ==
                                                  [(r.n_left, r.n_right, r.n_matched) for r in res]}
    try:
        dig = sequence.sequence_digest(merged, n_pairs)
        out["digest"] = dig
        gold = json.loads((ROOT / "tests" / "golden" / "kseq_digests.json").read_text())
        want = gold["prefix"].get(str(n_pairs))
        out["digest_equals_reference"] = (dig == want) if want else None
        out["digest_reference"] = "tests/golden/kseq_digests.json: the reference's own src/Frame.cc + ORBextractor.cc (oracle/_ref/libreframe.so) over the same pairs"
    except Exception as e:  # noqa: BLE001
        out["digest"] = f"unavailable ({type(e).__name__}: {e})"
    return out


def kseq_cpu_arm(cores: int) -> dict:
    """The reference's stereo Frame constructor (libreframe.so: two extraction threads + ComputeStereoMatches + grid) on
    all host cores, one process per two cores, over a bounded sample of the same pairs."""
    procs = max(1, cores // 2)
    runs = []
    for _ in range(3):
        out = subprocess.run([sys.executable, "-m", "oracle.cpu_bench", "--stereo", "--procs", str(procs), "--frames-per-proc", "12"],
                             cwd=str(ROOT), capture_output=True, text=True, check=True)
        runs.append(json.loads(out.stdout.strip().splitlines()[-1]))
    v = sorted(r["frames_per_s"] for r in runs)
    return {"pairs_per_s": v[1], "min": v[0], "max": v[2], "cores": cores, "kind": "reference",
            "sample": f"{runs[0]['frames']} pairs per run ({procs} processes x 12 pairs, 2 threads each as src/Frame.cc:110-113), median of 3 runs"}


def run_ours(args, rank, world, local):
    import torch
    from orb_slam2_test_b200 import ORBextractor, OrbStream, synth
    from orb_slam2_test_b200.extractor import pinned_empty
    from orb_slam2_test_b200.sequence import shard_indices

    torch.cuda.set_device(local)
    all_cpus = os.sched_getaffinity(0)
    numa_note = bind_near_gpu(local)  # before any pinned allocation (first touch)
    F = args.frames_per_step
    W = max(args.warmup, 3)
    # this rank's shard of the synthetic sequence: frame f belongs to rank f mod world (frames are a
    # pure function of their index); a step's batch cycles through the shard's first UNIQUE frames
    mine = list(shard_indices(UNIQUE * world, rank, world))
    uniq = np.stack([synth.make_frame((ROWS, COLS), seed=7, frame=f) for f in mine])
    batch = np.ascontiguousarray(uniq[np.arange(F) % UNIQUE])
    dev = torch.from_numpy(batch).cuda(local)
    torch.cuda.synchronize()

    ext = ORBextractor(NFEATURES, SCALE, NLEVELS, INI_TH, MIN_TH, device=local, max_batch=F,
                       want_pyramid=False)
    launches_per_call = ext.launches_per_call()
    # stage events are recorded INSIDE the timed calls (event-record nodes of the captured graph)
    stage_on = os.environ.get("BENCH_NO_STAGE_TIMING") != "1"  # A/B runs of options that stage timing disables
    ext.set_stage_timing(stage_on)

    # ---- value: device-resident inputs, CUDA events on the library's stream ----
    for _ in range(W):
        ext.extract_device(dev.data_ptr(), F, COLS, ROWS)
    ext.sync()
    sampler = ClockSampler(local)
    barrier(world)
    sampler.start()
    ext.sync()
    stage_sum = {k: 0.0 for k in ext.STAGES}
    ext.timer_start()
    for _ in range(args.steps):
        ext.extract_device(dev.data_ptr(), F, COLS, ROWS)
    ext.timer_stop()
    ext.sync()
    ms = ext.timer_ms()
    last_stage = ext.last_stage_ms() if stage_on else {}  # of the last timed call
    clocks = sampler.stop()
    barrier(world)
    ms_max = all_max(ms, world)
    value = world * F * args.steps / (ms_max / 1e3)

    # per-stage averages over K more calls of the same graph (each read back right after its call)
    for _ in range(args.steps):
        ext.extract_device(dev.data_ptr(), F, COLS, ROWS)
        st1 = ext.last_stage_ms() if stage_on else ext.stage_times(dev.data_ptr(), F, COLS, ROWS)
        for k in stage_sum:
            stage_sum[k] += st1[k]
    st = {k: v / args.steps for k, v in stage_sum.items()}

    # parity spot check of what was just timed (frame 0 of the batch) + mean keypoint count
    res = ext.fetch()
    n_kp = float(np.mean([len(k) for k, _ in res[:UNIQUE]]))
    n_cand = float(np.mean([sum(len(ext.debug_candidates(f, l)[0]) for l in range(NLEVELS)) for f in range(min(4, F))]))
    parity = "skipped"
    if rank == 0:
        try:
            from oracle.pyoracle import Oracle
            ok, od = Oracle(NFEATURES, SCALE, NLEVELS, INI_TH, MIN_TH).extract(batch[0])
            gk, gd = res[0]
            parity = "bit-exact" if (gk.tobytes() == ok.tobytes() and gd.tobytes() == od.tobytes()) else "MISMATCH"
        except Exception as e:  # the checker is optional here; tests/ are the parity gate
            parity = f"unavailable ({type(e).__name__})"

    if args.quick:
        if rank == 0:
            print(json.dumps({"quick": True, "value": value, "ms_per_step": ms_max / args.steps, "stage_ms": st,
                              "parity_check": parity, "keypoints_per_frame": n_kp, "candidates_per_frame": n_cand}),
                  flush=True)
        return
    ext.set_stage_timing(False)

    # ---- e2e: the library's pipelined feeder (borb_stream_*), page-locked host buffers, H2D of every frame and
    #      D2H of every result inside the timed region ----
    ext.close()
    sub = min(int(os.environ.get("BENCH_E2E_SUB", "128")), F)  # frames per push: small launches lose throughput to tails
    n_sub = max(1, F // sub)
    n_streams = max(1, int(os.environ.get("BENCH_E2E_STREAMS", "2")))
    slots = max(2, int(os.environ.get("BENCH_E2E_SLOTS", "3")))
    host = pinned_empty((n_sub * sub, ROWS, COLS))
    host[:] = batch[np.arange(n_sub * sub) % F]

    def e2e_measure(want_pyramid: bool, steps: int, warm: int) -> dict:
        """steps x n_sub pushes of `sub` frames as ONE continuous pipeline over n_streams rings (drained only at
        the end, inside the timing); wall clock bracketed by barriers, max over ranks."""
        streams = [OrbStream(NFEATURES, SCALE, NLEVELS, INI_TH, MIN_TH, COLS, ROWS, batch=sub, slots=slots, device=local,
                             want_pyramid=want_pyramid) for _ in range(n_streams)]
        fsz = ROWS * COLS
        seen = {"kp": 0, "batches": 0}

        def drain(st):
            _, m = st.pop()
            seen["kp"] += sum(st.count(j) for j in range(0, m, 16)) * 16  # every 16th item's count, cheaply
            seen["batches"] += 1
            st.release()

        def run(n_steps):
            seen["kp"] = seen["batches"] = 0
            for i in range(n_steps * n_sub):
                st = streams[i % n_streams]
                if st.slots_free() == 0:
                    drain(st)
                st.push_ptr(host.ctypes.data + (i % n_sub) * sub * fsz, None, sub, tag=i)
            for st in streams:
                while st.in_flight():
                    drain(st)

        run(warm)
        barrier(world)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        run(steps)
        torch.cuda.synchronize()
        ms_ = (time.perf_counter() - t0) * 1e3
        barrier(world)
        ms_all = all_max(ms_, world)
        up_b, down_b = streams[0].link_bytes()
        launches = streams[0].launches_per_push() * steps * n_sub
        for st in streams:
            st.close()
        frames = n_sub * sub * steps
        return {"value": world * frames / (ms_all / 1e3), "ms": ms_all, "h2d_bytes_per_step": n_sub * up_b,
                "d2h_bytes_per_step": n_sub * down_b, "launches": launches, "batches": seen["batches"],
                "GBps": world * frames / sub * (up_b + down_b) / (ms_all / 1e3) / 1e9}

    e2e = e2e_measure(False, args.steps, W)
    e2e_value = e2e["value"]
    # the same loop with every frame's padded pyramid brought back too (mvImagePyramid for a host-side consumer such
    # as the reference's own ComputeStereoMatches, src/Frame.cc:626,738,751,757): SURVEY.md 8 a11 "reported both ways"
    e2e_pyr = e2e_measure(True, max(2, args.steps // 4), 1)
    # raw copy ceiling of the box at this N, same sizes, no kernels (orb_slam2_test_b200/hostlink.py)
    from orb_slam2_test_b200 import hostlink
    link = hostlink.measure(local, sub * ROWS * COLS, e2e["d2h_bytes_per_step"] // n_sub, iters=40,
                            barrier=(lambda: barrier(world)) if world > 1 else None)
    link_s = all_max(link["both_s"], world)
    link_GBps = world * (link["h2d_bytes"] + link["d2h_bytes"]) * link["iters"] / link_s / 1e9
    link_up_s = all_max(link["h2d_alone_s"], world)
    link_up_GBps = world * link["h2d_bytes"] * link["iters"] / link_up_s / 1e9
    del host

    kseq = None
    if not os.environ.get("BENCH_NO_KSEQ"):
        try:
            kseq = kseq_section(args, rank, world, local, synth, OrbStream, pinned_empty)
        except Exception as e:  # noqa: BLE001 -- an extra key must not take the headline line down
            kseq = {"error": f"{type(e).__name__}: {e}"}
            barrier(world)

    # ---- single-frame latency through the host ABI (pageable input, like a cv::Mat) ----
    lat = {}
    stereo = prestep = poststep = matcher = tum = bow = None
    if rank == 0:
        one = ORBextractor(NFEATURES, SCALE, NLEVELS, INI_TH, MIN_TH, device=local, max_batch=1, want_pyramid=False)
        img = batch[0].copy()
        for want_pyr, name in ((False, "p50_us"), (True, "p50_us_with_pyramid")):
            for _ in range(20):
                one._lib.borb_extract_host(one._h, img.ctypes.data, 1, COLS, ROWS, COLS, 0, 1 if want_pyr else 0)
            ts = []
            for _ in range(300):
                t1 = time.perf_counter()
                one._lib.borb_extract_host(one._h, img.ctypes.data, 1, COLS, ROWS, COLS, 0, 1 if want_pyr else 0)
                ts.append((time.perf_counter() - t1) * 1e6)
            lat[name] = float(np.percentile(ts, 50))
            lat[name.replace("p50", "p99")] = float(np.percentile(ts, 99))
        # the same call with the frame in page-locked memory (borb_host_alloc): uploaded from where it lies, no pack
        from orb_slam2_test_b200.extractor import pinned_empty
        pimg = pinned_empty(img.shape)
        pimg[...] = img
        for _ in range(20):
            one._lib.borb_extract_host(one._h, pimg.ctypes.data, 1, COLS, ROWS, COLS, 0, 0)
        ts = []
        for _ in range(300):
            t1 = time.perf_counter()
            one._lib.borb_extract_host(one._h, pimg.ctypes.data, 1, COLS, ROWS, COLS, 0, 0)
            ts.append((time.perf_counter() - t1) * 1e6)
        lat["p50_us_page_locked_input"] = float(np.percentile(ts, 50))
        dptr = dev.data_ptr()
        ks = []
        for _ in range(200):
            one.extract_device(dptr, 1, COLS, ROWS)
            one.sync()
            ks.append(one.last_device_ms() * 1e3)
        lat["kernels_only_p50_us"] = float(np.percentile(ks, 50))
        # BASELINE config 4: a stereo pair = two extractor instances driven concurrently from two
        # host threads on two streams, as src/Frame.cc:110-113 does
        two = [one, ORBextractor(NFEATURES, SCALE, NLEVELS, INI_TH, MIN_TH, device=local, max_batch=1, want_pyramid=True)]
        imgs = [img, uniq[1 % len(uniq)].copy()]

        def eye(i):
            two[i]._lib.borb_extract_host(two[i]._h, imgs[i].ctypes.data, 1, COLS, ROWS, COLS, 0, 1)

        ts = []
        for it in range(220):
            t1 = time.perf_counter()
            th = threading.Thread(target=eye, args=(1,))
            th.start()
            eye(0)
            th.join()
            if it >= 20:
                ts.append((time.perf_counter() - t1) * 1e6)
        lat["stereo_pair_p50_us_with_pyramids"] = float(np.percentile(ts, 50))
        stereo = stereo_section(ORBextractor, synth, local, two, dev)
        prestep = prestep_section(ORBextractor, synth, local, measured_peak()[0])
        poststep = poststep_section(ORBextractor, synth, local)
        matcher = matcher_section(ORBextractor, synth, local)
        try:
            bow = bow_section(ORBextractor, synth, local)
        except Exception as ex:  # a reported side measurement, never the headline
            bow = {"unavailable": f"{type(ex).__name__}: {ex}"}
        tum = tum_section(ORBextractor, synth, local)

    if rank != 0:
        return
    peak, peak_src = measured_peak()
    stage_bytes = stage_algorithmic_bytes(n_kp, n_cand)
    dom = max(st, key=st.get)
    dom_s = st[dom] / 1e3
    achieved = stage_bytes[dom] * F / dom_s / 1e9
    step_ms = ms_max / args.steps
    os.sched_setaffinity(0, all_cpus)  # the CPU baseline uses every host core
    cores = len(os.sched_getaffinity(0))
    cpu = None
    if world == 1:
        cpu = cpu_baseline_stable(cores)
        r1 = cpu_reference_run(1, 16)
        cpu["single_core_frames_per_s"] = r1["frames_per_s"]
        try:
            out = subprocess.run([sys.executable, "-m", "oracle.cpu_bench", "--primitives"], cwd=str(ROOT), capture_output=True,
                                 text=True, check=True)
            cpu["primitive_ratio"] = json.loads(out.stdout.strip().splitlines()[-1])
        except Exception as e:  # noqa: BLE001
            cpu["primitive_ratio"] = {"unavailable": type(e).__name__}
        if isinstance(kseq, dict) and "pairs_per_s" in kseq:
            try:
                kseq["cpu_baseline"] = kseq_cpu_arm(cores)
            except Exception as e:  # noqa: BLE001
                kseq["cpu_baseline"] = {"unavailable": type(e).__name__}
    line = {
        "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": W, "ms_per_step": step_ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": bench_config(F),
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "frames/s",
                "h2d_bytes_per_step": e2e["h2d_bytes_per_step"],
                "d2h_bytes_per_step": e2e["d2h_bytes_per_step"],
                "link_GBps": e2e["GBps"],
                "link_ceiling_GBps": link_GBps, "frac_of_link": e2e["GBps"] / link_GBps,
                "link_ceiling_h2d_alone_GBps": link_up_GBps,
                "frames_per_s_link_ceiling": world * sub * link["iters"] / link_s,
                "gpu_launches": e2e["launches"],
                "with_pyramid": {"value": e2e_pyr["value"], "unit": "frames/s", "h2d_bytes_per_step": e2e_pyr["h2d_bytes_per_step"],
                                 "d2h_bytes_per_step": e2e_pyr["d2h_bytes_per_step"], "link_GBps": e2e_pyr["GBps"],
                                 "what": "same loop with BORB_STREAM_PYRAMID: every frame's 8 padded levels (1.79 MB) come back too, "
                                         "what a stock Frame.cc caller of mvImagePyramid pays"},
                "how": f"{n_streams} library feeder rings (borb_stream_*) x {slots} slots x {sub}-frame batches: upload, kernel and "
                       f"download streams per ring, page-locked host memory uploaded in place, wall clock between barriers, max over "
                       f"ranks; link ceiling = plain cudaMemcpyAsync of the same sizes on every rank at once on two upload + two download streams (as the two rings), best of 3 runs, no kernels; {numa_note}"},
        "gpu_launches": args.steps * launches_per_call,
        "roofline": {"bound": "hbm", "kernel": STAGE_KERNEL[dom], "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": ncu_traffic(STAGE_KERNEL[dom], F), "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": stage_bytes[dom] * F,
                     "kernel_ms_per_launch": st[dom], "kernel_share_of_step": st[dom] / step_ms,
                     "how": "CUDA events recorded between the stages inside the timed graph launches "
                            "(borb_set_stage_timing), averaged over the timed calls",
                     "stage_ms_per_launch": st, "stage_ms_last_timed_call": last_stage,
                     "stage_GBps": {k: stage_bytes[k] * F / (st[k] / 1e3) / 1e9 for k in st},
                     "path_achieved": algorithmic_bytes(n_kp) * value / world / 1e9,
                     "path_frac": algorithmic_bytes(n_kp) * value / world / 1e9 / peak,
                     **issue_roofline(STAGE_KERNEL[dom], F, st[dom], clocks.get("sm_mhz") if isinstance(clocks, dict) else None)},
        "cpu_baseline": cpu,
        "kseq": kseq,
        "latency": lat,
        "stereo": stereo,
        "prestep": prestep,
        "poststep": poststep,
        "matcher": matcher,
        "bow": bow,
        "tum": tum,
        "keypoints_per_frame": n_kp, "candidates_per_frame": n_cand,
        "parity_check": parity,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--frames-per-step", type=int, default=512)
    ap.add_argument("--quick", action="store_true", help="value loop + stage times only (for ncu runs)")
    args = ap.parse_args()
    if args.impl == "reference":
        rank = int(os.environ.get("RANK", "0"))
        run_reference_arm(args, rank, int(os.environ.get("WORLD_SIZE", "1")))
        return
    rank, world, local = dist_setup(args.gpus)
    try:
        run_ours(args, rank, world, local)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
