#!/usr/bin/env python
"""Headline benchmark: swapped 256x256 frames per second (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W                       (our arm; under torchrun for N > 1)
  python bench.py --impl reference --gpus N --steps K --warmup W      (the reference's CPU path, host cores)

One step = one batch of B = 64 faces per GPU through the whole hot path (BASELINE.json configs[1] plus the
paste-back of configs[2]):  ArcFace iresnet100 z_id of 64 synthetic 224x224 sources  ->  AEI_Net('unet', 2, 512)
on 64 synthetic aligned 256x256 crops  ->  fused uint8 conversion + 256->224 resize + inverse warp + blurred-mask
composite into 64 synthetic 1080p frames (one face per frame).  Weights are random-init of the named architectures
(no network access), data is synthetic.  Frames shard across GPUs with no collective inside the step (weak scaling).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.dont_write_bytecode = True

import numpy as np  # noqa: E402
import torch  # noqa: E402

METRIC = "swapped_256x256_frames_per_sec"
UNIT = "frames/s"
B = 64
H, W = 1080, 1920
BACKBONE, NUM_BLOCKS = "unet", 2
# 13.2 MB per 1080p frame*face = frame read + write 12.44 MB, + per face Y 0.39 MB + mask <= 0.4 MB
PASTE_BYTES_PER_FRAME, PASTE_BYTES_PER_CROP = 12.44e6, 0.79e6


_REAL_STDOUT = None


def quiet_stdout():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner on fd 1),
    so everything else is sent to stderr for the duration of the run and the line goes to the saved descriptor."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_REAL_STDOUT, data)


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(hbm=p["hbm_gbs"], tf_burst=p["bf16_tflops"], tf_sustained=p["bf16_tflops_sustained"],
                    source="measured (MEASURED_PEAKS.json)")
    except Exception:
        return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback (B200_PROFILING.md)")


class ClockSampler(threading.Thread):
    """SM clock / throttle reasons DURING the timed region: NVML polled every 10 ms (the timed region of the default
    run is ~0.2 s), nvidia-smi every 0.2 s where NVML cannot be loaded."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    BITS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], threading.Event()
        self.nv = self.handle = None
        try:
            import pynvml
            pynvml.nvmlInit()
            try:
                uuid = str(torch.cuda.get_device_properties(index).uuid)
                self.handle = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
            except Exception:
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nv = pynvml
        except Exception:
            self.nv = None

    def run(self):
        while not self.stop_flag.is_set():
            try:
                if self.nv is not None:
                    nv, h = self.nv, self.handle
                    mhz = float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                    try:
                        mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))
                    except Exception:
                        mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                    watts = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
                    self.rows.append([mhz, self.max_mhz, watts] + [bool(mask & b) for _, b in self.BITS])
                else:
                    out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                    parts = [x.strip() for x in out.strip().split(",")]
                    if len(parts) >= 7:
                        self.rows.append([float(parts[0]), float(parts[1]), float(parts[2])] +
                                         [p.lower().startswith("active") for p in parts[3:7]])
            except Exception:
                pass
            self.stop_flag.wait(0.01 if self.nv is not None else 0.2)

    def summary(self):
        self.stop_flag.set()
        self.join(timeout=6)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no clock samples (NVML and nvidia-smi unavailable)"]}
        sm = sorted(r[0] for r in self.rows)
        reasons = [n for i, (n, _) in enumerate(self.BITS) if any(r[3 + i] for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.rows[0][1], "sm_min_mhz": sm[0], "reasons": reasons,
                "power_w_max": max(r[2] for r in self.rows), "samples": len(self.rows),
                "source": "nvml, 10 ms period" if self.nv is not None else "nvidia-smi, 0.2 s period"}


WORKLOADS = {
    # BASELINE.json configs[1] + the paste-back of configs[2]: one face per frame, one source per face
    "default": dict(faces=64, frames=64, sources=64, faces_per_frame=1,
                    text="configs[1] + paste-back: AEI_Net(unet, num_blocks=2, c_id=512) fp16 (the reference's G.half()) + "
                         "ArcFace iresnet100 z_id, batch 64 aligned 256x256 synthetic crops per GPU, fused paste-back "
                         "into 64 1080p frames (1 face/frame)"),
    # BASELINE.json configs[3]: 4 source/target pairs per frame - 4 sources embedded once per step, 16 frames x 4
    # crops per step, the four faces of a frame composited in list order (two frames in 16 have overlapping boxes)
    "config4": dict(faces=64, frames=16, sources=4, faces_per_frame=4,
                    text="configs[3]: multi-face video, 4 source/target pairs per frame: ArcFace z_id of 4 sources + "
                         "AEI_Net(unet, num_blocks=2, c_id=512) fp16 on 64 crops (16 frames x 4 faces) per GPU + fused "
                         "paste-back of 4 faces per 1080p frame (overlapping boxes in 2 of 16 frames)"),
}


def make_inputs(seed, wl="default"):
    """Synthetic batch (numpy, host): sources, crops, frames, masks, affines, face -> frame / source maps."""
    from ghost_b200 import synth
    w = WORKLOADS[wl]
    nf, nF, nS, per = w["faces"], w["frames"], w["sources"], w["faces_per_frame"]
    rng = np.random.default_rng(100 + seed)
    sources = synth.sources_u8(nS, seed=seed)
    crops = synth.crops_u8(nf, seed=seed)
    base = synth.frame_u8(H, W, seed=seed)
    frames = np.stack([np.roll(base, 17 * i, axis=1) for i in range(nF)])
    masks = np.stack([synth.soft_mask(rng) for _ in range(nf)])
    if per == 1:
        tfms = np.stack([synth.similarity_tfm(H, W, rng, clipped=(i % 20 == 19)) for i in range(nf)])
        face_frame = np.arange(nf, dtype=np.int32)
        face_source = np.arange(nf, dtype=np.int64)
    else:
        tfms = []
        for f in range(nF):
            for q in range(per):
                # one face per quadrant; in every eighth frame face 1 sits on top of face 0 (overlapping boxes)
                qx, qy = (q % 2), (q // 2)
                if f % 8 == 7 and q == 1:
                    qx, qy = 0, 0
                cx = W * (0.25 + 0.5 * qx) + float(rng.uniform(-60, 60))
                cy = H * (0.25 + 0.5 * qy) + float(rng.uniform(-30, 30))
                tfms.append(synth.similarity_tfm(H, W, rng, scale=float(rng.uniform(0.7, 0.95)), center=(cx, cy)))
        tfms = np.stack(tfms)
        face_frame = np.repeat(np.arange(nF, dtype=np.int32), per)
        face_source = np.tile(np.arange(per, dtype=np.int64), nF) % nS
    return dict(sources=sources, crops=crops, frames=frames, masks=masks, tfms=tfms, face_frame=face_frame,
                face_source=face_source)


def make_arcface(tdt):
    """iresnet100, random init of the named architecture, `features` running_var set so that |z_id| ~ 22 like a
    trained ArcFace (weights.unit_embedding_); the CPU baseline and the eager-GPU leg use the same state dict."""
    import ghost_b200
    from ghost_b200 import weights as Wt
    arc = ghost_b200.iresnet100().eval()
    arc.load_state_dict(Wt.unit_embedding_(arc.state_dict()))
    return arc.to(tdt).cuda()


def cpu_baseline_run(wl, n_faces, steps, warmup):
    """The reference's CPU path (oracle port: same torch ops / cv2 calls) on all host cores, on the first n_faces
    faces (whole frames) of the workload's step.  Returns (frames/s, s/step, cores, frames of the last step, swapped
    crops of the last step)."""
    from ghost_b200 import weights as Wt
    from oracle import pipeline_cpu
    w = WORKLOADS[wl]
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    aei_sd = Wt.init_state_dict(Wt.aei_spec(BACKBONE, NUM_BLOCKS, 512), seed=0)
    arc_sd = Wt.unit_embedding_(Wt.init_state_dict(Wt.iresnet100_spec(), seed=0))
    d = make_inputs(0, wl)
    per = w["faces_per_frame"]
    n_frames = max(1, n_faces // per)
    n_faces = n_frames * per
    sl = slice(0, n_faces)
    n_src = min(w["sources"], n_faces) if per == 1 else w["sources"]
    args = (aei_sd, arc_sd, BACKBONE, NUM_BLOCKS, d["sources"][:n_src], d["crops"][sl], d["frames"][:n_frames],
            d["masks"][sl], d["tfms"][sl])
    kw = dict(face_frame=d["face_frame"][sl], face_source=d["face_source"][sl], return_swapped=True)
    for _ in range(warmup):
        pipeline_cpu.swap_frames_cpu(*args, **kw)
    t0 = time.perf_counter()
    for _ in range(steps):
        out, swapped = pipeline_cpu.swap_frames_cpu(*args, **kw)
    dt = time.perf_counter() - t0
    return n_faces * steps / dt, dt / steps, cores, out, swapped, n_faces


def run_reference(args):
    """--impl reference: the reference's CPU path for the same workload on all host cores (rank 0 only).  Every step
    is a bounded sample of the workload's step (4 faces of the 64), --steps / --warmup are honoured."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    if args.workload == "config5":
        return run_reference_config5(args)
    wl = args.workload
    steps, warm = max(1, args.steps), max(0, args.warmup)
    fps, sec_per_step, cores, _, _, n_faces = cpu_baseline_run(wl, 4, steps, warm)
    sample = ("%d of the %d faces of one step per timed step (ArcFace fp32 + AEI_Net fp32 + cv2 paste-back into 1080p), "
              "%d timed steps, %d warm-up" % (n_faces, WORKLOADS[wl]["faces"], steps, warm))
    line = {
        "impl": "reference", "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": sec_per_step * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.gpus, wl),
        "cpu_baseline": {"value": fps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


def workload_config(n_gpus, wl="default"):
    w = WORKLOADS[wl]
    return {"workload": w["text"],
            "faces_per_step_per_gpu": w["faces"], "frames_per_step_per_gpu": w["frames"], "frame": "%dx%d" % (W, H),
            "backbone": BACKBONE, "num_blocks": NUM_BLOCKS,
            "parallelism": "frames sharded over %d GPU(s), no collective in the step" % n_gpus,
            "l2": "inputs larger than L2: per-step working set (weights ~0.5 GB + frames + activations ~6 GB) "
                  "exceeds the 126 MB L2, no flush needed"}


def eager_gpu_reference(dev, d, wl):
    """The reference's own deployment on the SAME GPU: PyTorch eager with cuDNN / cuBLAS,
    `G.cuda().half()` + `netArc.cuda()` in fp32 (inference.py:26-36), cudnn.benchmark on.  /root/reference does not
    travel to the GPU box, so the modules are the oracle's torch-functional restatements (oracle/aei_ref.py,
    oracle/iresnet_ref.py: the same torch ops in the same order) moved to the device.  Not the product path."""
    import torch.nn.functional as F
    from ghost_b200 import weights as Wt
    from oracle import aei_ref, iresnet_ref
    w = WORKLOADS[wl]
    keep = torch.backends.cudnn.benchmark
    torch.backends.cudnn.benchmark = True
    try:
        with torch.no_grad():
            aei_sd = {k: v.to(dev).half() for k, v in Wt.init_state_dict(Wt.aei_spec(BACKBONE, NUM_BLOCKS, 512), seed=0).items()
                      if v.dtype.is_floating_point}
            arc_sd = {k: v.to(dev) for k, v in Wt.unit_embedding_(Wt.init_state_dict(Wt.iresnet100_spec(), seed=0)).items()
                      if v.dtype.is_floating_point}
            src = torch.from_numpy(d["sources"]).to(dev)
            crp = torch.from_numpy(d["crops"]).to(dev)
            fsrc = torch.from_numpy(d["face_source"]).to(dev)

            def arcface():
                x = src.float() / 255.                                         # normalize_and_torch_batch
                x = (x.permute(0, 3, 1, 2) - 0.5) / 0.5
                x = F.interpolate(x, scale_factor=0.5, mode="bilinear", align_corners=True)
                return iresnet_ref.iresnet100_forward(arc_sd, x)

            def generator(z):
                t = crp[:, :, :, [2, 1, 0]] / 255.                             # transform_target_to_torch(half=True)
                t = ((t.half() - 0.5) / 0.5).permute(0, 3, 1, 2)
                attrs = aei_ref.encode(aei_sd, t, BACKBONE)
                Y = aei_ref.generate(aei_sd, attrs, z.half()[fsrc], NUM_BLOCKS)
                Y = (Y.permute(0, 2, 3, 1) * 0.5 + 0.5) * 255                  # faceshifter_batch
                return Y[:, :, :, [2, 1, 0]].type(torch.uint8)

            for _ in range(3):
                z = arcface()
                generator(z)
            torch.cuda.synchronize()
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            reps = 5
            t_arc = t_gen = 0.0
            for _ in range(reps):
                ev[0].record()
                z = arcface()
                ev[1].record()
                generator(z)
                ev[2].record()
                torch.cuda.synchronize()
                t_arc += ev[0].elapsed_time(ev[1])
                t_gen += ev[1].elapsed_time(ev[2])
            t_arc, t_gen = t_arc / reps, t_gen / reps
        return {"arcface_ms": t_arc, "generator_ms": t_gen,
                "value": w["faces"] / ((t_arc + t_gen) * 1e-3), "unit": "faces/s (ArcFace + generator, no paste-back)",
                "kind": "port on cuda: oracle/aei_ref.py + oracle/iresnet_ref.py (the reference's torch ops) as "
                        "G.half() + netArc fp32, cuDNN/cuBLAS eager, cudnn.benchmark=True, %d crops, %d sources" % (
                            w["faces"], w["sources"])}
    finally:
        torch.backends.cudnn.benchmark = keep
        torch.cuda.empty_cache()


def run_ours(args):
    import torch.distributed as dist
    import ghost_b200

    wl = args.workload
    w = WORKLOADS[wl]
    B, NF, NS = w["faces"], w["frames"], w["sources"]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device (there is no CPU fallback)")
    # the CPU baseline runs first, alone on the host cores (N = 1 only: at N > 1 the other ranks would spin in a
    # barrier beside it): 2 whole frames of the step, 1 warm-up + 2 timed passes
    cpu = None
    if world == 1:
        cpu = cpu_baseline_run(wl, 2 * w["faces_per_frame"], 2, 1)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # one process per GPU: sit on the CPUs / memory of the GPU's NUMA node before any pinned buffer is allocated
    from ghost_b200.shard import bind_to_gpu_numa
    numa = bind_to_gpu_numa(local) if os.environ.get("GHOST_NUMA_BIND", "1") != "0" else None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    pk = peaks()
    # everything runs on an explicit stream: the legacy default stream serialises against every other blocking
    # stream, which would undo the overlap of the copy / encoder side streams
    torch.cuda.set_stream(torch.cuda.Stream(device=dev))

    # .half() = fp16 operands on tcgen05 kind::f16: the reference's deployment dtype (inference.py:30) and the mode
    # that meets north_star's 1e-2 / 1e-3 tolerance; --dtype bf16 times the bf16-storage variant of the same kernels
    tdt = torch.float16 if args.dtype == "fp16" else torch.bfloat16
    G = ghost_b200.AEI_Net(BACKBONE, NUM_BLOCKS, 512, seed=0).eval().to(tdt).cuda()
    arc = make_arcface(tdt)
    G.enable_graph(True)
    arc.enable_graph(True)

    d = make_inputs(rank, wl)
    tfms, face_frame = d["tfms"], d["face_frame"]
    pin = lambda a: torch.from_numpy(a).pin_memory()
    sources_p, crops_p, frames_p = pin(d["sources"]), pin(d["crops"]), pin(d["frames"])
    out_p = torch.empty_like(frames_p).pin_memory()
    masks_p = pin(d["masks"])
    masks_d = masks_p.to(dev)
    sources_d, crops_d, frames_d = sources_p.to(dev), crops_p.to(dev), frames_p.to(dev)
    out_d = torch.empty_like(frames_d)
    fsrc_d = torch.from_numpy(d["face_source"]).to(dev) if NS != B else None

    side, side_ev = torch.cuda.Stream(device=dev), torch.cuda.Event()
    copy_st, copy_ev = torch.cuda.Stream(device=dev), torch.cuda.Event()

    def step_device(src, crp, frm, out):
        # The frames of the step go into the output buffer first - a device-to-device copy that depends on nothing the
        # networks compute, so it runs on its own stream while they work (out-of-place paste-back = this copy + the
        # in-place composite below; paste_back(frames, ..., out=other) does the same copy on the calling stream).
        main = torch.cuda.current_stream(dev)
        copy_st.wait_stream(main)                    # the previous step is done with `out`
        with torch.cuda.stream(copy_st):
            out.copy_(frm, non_blocking=True)
            copy_ev.record(copy_st)
        # ArcFace on the current stream, the attribute encoder (independent of z_id) beside it on `side`, then the
        # AAD generator; the generator output stays on the device
        ghost_b200.embed_and_swap(G, arc, src, crp, side=side, event=side_ev, write_u8=False, face_source=fsrc_d)
        yptr, ydt = G.y_nhwc(B, dev)
        main.wait_event(copy_ev)
        ghost_b200.paste_back(out, (yptr, ydt, B), masks_d, tfms, face_frame, mode="exact", out=out)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------------------------------------------------------- device-resident timing (value)
    for _ in range(max(args.warmup, 3)):
        step_device(sources_d, crops_d, frames_d, out_d)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step_device(sources_d, crops_d, frames_d, out_d)
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    clocks = sampler.summary()

    # ---------------------------------------------------------------- parity of the timed step (rank 0, N = 1)
    parity = None
    if cpu is not None:
        from oracle import pipeline_cpu
        ref_frames, ref_swapped, n_cmp = cpu[3], cpu[4], cpu[5]
        nfr = ref_frames.shape[0]
        got = out_d[:nfr].cpu().numpy()
        dfr = np.abs(got.astype(np.int32) - ref_frames.astype(np.int32))
        # the generator output as uint8 crops (same launches, uint8 conversion on), then the reference's cv2
        # paste-back of exactly those crops on the host: must reproduce the device frames bit for bit
        u8 = ghost_b200.embed_and_swap(G, arc, sources_d, crops_d, write_u8=True, face_source=fsrc_d)
        torch.cuda.synchronize()
        u8h = u8[:n_cmp].cpu().numpy()
        dsw = np.abs(u8h.astype(np.int32) - ref_swapped.astype(np.int32))
        exact = True
        for f in range(nfr):
            sel = [i for i in range(n_cmp) if face_frame[i] == f]
            cvf = pipeline_cpu.paste_cv2(d["frames"][f], [u8h[i] for i in sel], [d["masks"][i] for i in sel],
                                         [tfms[i] for i in sel])
            exact = exact and bool(np.array_equal(cvf, got[f]))
        parity = {"faces": int(n_cmp), "frames": int(nfr),
                  "max_abs_u8": int(dfr.max()), "mean_abs_u8": float(dfr.mean()),
                  "swapped_crops_max_abs_u8": int(dsw.max()), "swapped_crops_mean_abs_u8": float(dsw.mean()),
                  "swapped_crops_frac_beyond_1_level": float((dsw > 1).mean()),
                  "paste_bit_exact_given_Y": exact,
                  "against": "the CPU port's fp32 forward + cv2 paste-back of the same faces (frames of the last "
                             "timed step); 1 grey level = 7.8e-3 on Y in [-1,1]"}

    # ---------------------------------------------------------------- end-to-end timing with host buffers (e2e)
    # The public streaming call: every step uploads its sources, crops, frames and masks from pinned host memory
    # and downloads the composited frames; SwapPipeline double-buffers the three stages over CUDA streams, the
    # last download is inside the timed region (flush before the closing event).
    def time_e2e(pipe, out):
        for _ in range(3):
            pipe.submit(sources_p, crops_p, frames_p, masks_p, tfms, face_frame, out, face_source=fsrc_d)
        pipe.flush()
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for _ in range(args.steps):
            pipe.submit(sources_p, crops_p, frames_p, masks_p, tfms, face_frame, out, face_source=fsrc_d)
        pipe.flush()
        f1.record()
        barrier()
        return f0.elapsed_time(f1)

    # (a) whole frames up, whole frames down into a second host buffer
    pipe_full = ghost_b200.SwapPipeline(G, arc, B, (H, W), depth=2, mode="exact", n_frames=NF, n_sources=NS)
    ms_e2e_full = time_e2e(pipe_full, out_p)
    h2d_full, d2h_full = pipe_full.h2d_bytes, pipe_full.d2h_bytes
    del pipe_full
    # (b) the frames stay in the caller's pinned buffer and are composited in place: only the face boxes (all the
    # paste-back touches) cross PCIe.  Same result in host memory; this is the e2e figure.  (The bench feeds the
    # same host frames to every step, so here the faces are composited onto the previous step's result.)
    pipe = ghost_b200.SwapPipeline(G, arc, B, (H, W), depth=2, mode="exact", boxes_only=True, n_frames=NF, n_sources=NS)
    ms_e2e = time_e2e(pipe, frames_p)
    h2d, d2h_bytes = pipe.h2d_bytes, pipe.d2h_bytes
    del pipe

    # ---------------------------------------------------------------- with the final frame gather, N > 1
    # The only exchange of the path: every rank's finished uint8 frames go to rank 0 in frame order
    # (ghost_b200.FrameSharder: the gather of step k travels on a side stream while step k+1 computes);
    # device-resident like `value`, one gather per step, the last one inside the timed region.
    ms_gather = 0.0
    ms_gather_nccl = 0.0
    if world > 1:
        outs = [out_d, torch.empty_like(out_d)]

        def time_gather(start, finish, slot=None):
            def one(i):
                # the consumer rank composites straight into its slot of the gather buffer (no local copy)
                dst_t = slot() if slot is not None else None
                dst_t = outs[i % 2] if dst_t is None else dst_t
                step_device(sources_d, crops_d, frames_d, dst_t)
                start(dst_t)
            for i in range(2):
                one(i)
            finish()
            barrier()
            g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            g0.record()
            for i in range(args.steps):
                one(i)
            finish()
            g1.record()
            barrier()
            return g0.elapsed_time(g1)

        # (a) peer-mapped buffers: one copy-engine transfer per rank and step into rank 0's IPC-exported buffer
        peer = ghost_b200.PeerGather(NF * world, tuple(out_d.shape[1:]), dst=0, device=dev)
        ms_gather = time_gather(peer.gather_async, peer.wait, peer.local_out)
        peer.close()
        del peer
        # (b) the same exchange as NCCL send / recv (a kernel on both ends), for comparison
        sharder = ghost_b200.FrameSharder(NF * world, rank, world)
        ms_gather_nccl = time_gather(lambda t: sharder.gather_async(t, dst=0), sharder.wait)

    t = torch.tensor([ms_total, ms_e2e, ms_gather, ms_e2e_full, ms_gather_nccl], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_e2e, ms_gather, ms_e2e_full, ms_gather_nccl = (float(t[i]) for i in range(5))

    # ---------------------------------------------------------------- per-kernel-class roofline (rank 0)
    line = None
    if rank == 0:
        prof_g = G.profile(B, reps=3)
        prof_a = arc.profile(NS, reps=3)
        # paste kernel alone
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        yptr, ydt = G.y_nhwc(B, dev)
        # queue the paste launches behind a long-running forward so that the events bracket GPU time only
        # (the call's host-side preparation would otherwise dominate a 0.2 ms kernel)
        G.swap_u8(crops_d, torch.zeros(1, 512, dtype=tdt, device=dev), write_u8=False)
        p0.record()
        for _ in range(3):
            ghost_b200.paste_back(frames_d, (yptr, ydt, B), masks_d, tfms, face_frame, mode="exact", out=out_d)
        p1.record()
        torch.cuda.synchronize()
        paste_ms = p0.elapsed_time(p1) / 3

        def agg(prof, cls):
            sel = [o for o in prof if o["cls"] in cls]
            return (sum(o["ms"] for o in sel), sum(o["flops"] for o in sel), sum(o["bytes"] for o in sel), len(sel))
        conv_ms, conv_fl, _, conv_n = agg(prof_g + prof_a, (0,))
        aad_ms, _, aad_by, aad_n = agg(prof_g, (1,))
        st_ms, _, st_by, st_n = agg(prof_g, (3,))
        up_ms, _, up_by, up_n = agg(prof_g, (4,))
        all_ms = sum(o["ms"] for o in prof_g + prof_a) + paste_ms
        conv_tf = conv_fl / (conv_ms * 1e-3) / 1e12 if conv_ms > 0 else 0.0
        roofline = {"kernel": "igemm_kernel / igemm2_kernel (tcgen05 implicit-GEMM conv, one CTA or cta_group::2 pairs; %d launches/step incl. the two 64->3 convs of AADBlk8 on tapn_kernel)" % conv_n, "bound": "tensor",
                    "achieved": conv_tf, "peak": pk["tf_sustained"], "unit": "TFLOP/s",
                    "frac": conv_tf / pk["tf_sustained"], "traffic": ncu_traffic(),
                    "peak_source": pk["source"] + ", sustained",
                    "frac_of_burst_peak": conv_tf / pk["tf_burst"],
                    "share_of_step": conv_ms / all_ms if all_ms > 0 else None}

        def hbm(ms, by, n, name):
            a = by / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
            return {"kernel": name, "launches_per_step": n, "bound": "hbm", "achieved": a, "peak": pk["hbm"],
                    "unit": "GB/s", "frac": a / pk["hbm"], "ms_per_step": ms}
        roofline_hbm = [hbm(aad_ms, aad_by, aad_n, "igemm_kernel AAD mode (gamma/beta GEMM + fused IN/mask/blend/ReLU)"),
                        hbm(st_ms, st_by, st_n, "in_stats_kernel"), hbm(up_ms, up_by, up_n, "upsample2x_kernel"),
                        hbm(paste_ms, PASTE_BYTES_PER_FRAME * NF + PASTE_BYTES_PER_CROP * B, 3, "paste-back: frame copy + resize_crops_kernel + paste_faces_kernel (exact mode)")]

        n_launch_g, flops_g = G.stats(B)
        n_launch_a, flops_a = arc.stats(NS)
        launches_per_step = n_launch_g + n_launch_a + 2 + 2   # + preproc, arcface preproc, crop resize, face paste (the frame copy is a memcpy)
        fps = B * world * args.steps / (ms_total * 1e-3)
        fps_e2e = B * world * args.steps / (ms_e2e * 1e-3)
        gen_ms = sum(o["ms"] for o in prof_g)
        arc_ms = sum(o["ms"] for o in prof_a)
        line = {
            "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f16" if tdt == torch.float16 else "bf16", "data": "synthetic", "config": workload_config(world, wl),
            "host_numa_node": numa,
            "e2e": {"value": fps_e2e, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h_bytes),
                    "ms_per_step": ms_e2e / args.steps,
                    "note": "SwapPipeline(boxes_only=True): frames composited in place in the caller's pinned host "
                            "buffer, only the face boxes cross PCIe (bytes are per GPU)"},
            "e2e_full_frames": {"value": B * world * args.steps / (ms_e2e_full * 1e-3), "unit": UNIT,
                                "h2d_bytes_per_step": int(h2d_full), "d2h_bytes_per_step": int(d2h_full),
                                "ms_per_step": ms_e2e_full / args.steps,
                                "note": "SwapPipeline(): whole frames up, whole frames down into a second host buffer"},
            "gpu_launches": int(launches_per_step * args.steps),
            "clocks": clocks,
            "roofline": roofline, "roofline_hbm": roofline_hbm,
            "cpu_baseline": ({"value": cpu[0], "unit": UNIT, "cores": cpu[2], "kind": "port",
                              "sample": "%d of the %d faces of one step (whole frames), 1 warm-up + 2 timed passes, all "
                                        "host cores (torch fp32 + cv2), run before the GPU work" % (cpu[5], B)}
                             if cpu is not None else
                             {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                              "sample": "measured at N=1 only (the other ranks would spin beside it)"}),
            "parity": parity,
            "breakdown_ms_per_step": {"generator": gen_ms, "arcface": arc_ms,
                                      "paste_back": paste_ms, "note": "un-graphed per-op CUDA-event timing"},
            "generator_plus_paste_only": {
                "value": B / ((gen_ms + paste_ms) * 1e-3), "unit": UNIT,
                "note": "the reference embeds each source ONCE per video (core.py:52-54), not per frame: a long video "
                        "pays only generator + paste-back per frame; the headline step re-embeds 64 sources per step"},
            "flops_per_step": flops_g + flops_a,
        }
        if world == 1 and not args.no_eager:
            try:
                eg = eager_gpu_reference(dev, d, wl)
                eg["ours_same_scope_ms"] = gen_ms + arc_ms
                eg["speedup_same_scope"] = (eg["arcface_ms"] + eg["generator_ms"]) / (gen_ms + arc_ms)
                eg["speedup_generator"] = eg["generator_ms"] / gen_ms
                eg["speedup_arcface"] = eg["arcface_ms"] / arc_ms
                line["reference_gpu_eager"] = eg
            except Exception as ex:  # the baseline leg must not take the bench line down with it
                line["reference_gpu_eager"] = {"unavailable": "%s: %s" % (type(ex).__name__, str(ex)[:200])}
        if world > 1:
            line["with_gather"] = {"value": B * world * args.steps / (ms_gather * 1e-3), "unit": UNIT,
                                   "ms_per_step": ms_gather / args.steps,
                                   "bytes_into_rank0_per_step": int((world - 1) * out_d.numel()),
                                   "note": "device-resident step + gather of every rank's finished frames on rank 0 "
                                           "(frame order): ghost_b200.PeerGather - copy-engine transfers into rank "
                                           "0's IPC-mapped buffer, step k's frames travel while step k+1 computes, "
                                           "the last gather is inside the timed region",
                                   "nccl_send_recv": {"value": B * world * args.steps / (ms_gather_nccl * 1e-3),
                                                      "ms_per_step": ms_gather_nccl / args.steps,
                                                      "note": "the same exchange through FrameSharder.gather_async "
                                                              "(NCCL send / recv kernels beside the compute)"}}
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        emit(line)
    return 0


def ncu_traffic():
    """roofline.traffic: dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the
    committed ncu --set full capture (profiles/ncu_traffic.json, written by tools/ncu_summary.py); null if absent."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return json.load(f)["igemm_kernel"]
    except Exception:
        return None


C5_FACES, C5_SRB = 60, 20


def config5_config(world):
    return {"workload": "configs[4]: AEI_Net(linknet, num_blocks=3) + ArcFace z_id -> pix2pix SR generator "
                        "(LIPSPADEGenerator, train-mode batch statistics, batches of 20, 256->256 as the reference) "
                        "-> exact paste-back into 1080p frames, 60 faces per step per GPU",
            "faces_per_step_per_gpu": C5_FACES, "sr_batch": C5_SRB, "frame": "%dx%d" % (W, H),
            "parallelism": "frames sharded over %d GPU(s) on SR-batch boundaries, no collective in the step" % world,
            "l2": "inputs larger than L2 (weights ~0.8 GB + frames 0.75 GB per step)"}


def cpu_baseline_config5(steps, warmup):
    """configs[4] on the host cores (oracle ports of linknet/3, iresnet100 and the SR generator + cv2 paste-back):
    ONE SR batch of 20 faces per step - a third of the GPU step (the SR batch is the unit the reference's DataLoader
    forms, its batch statistics make smaller samples a different computation)."""
    from ghost_b200 import weights as Wt
    from oracle import pipeline_cpu
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    aei_sd = Wt.init_state_dict(Wt.aei_spec("linknet", 3, 512), seed=0)
    arc_sd = Wt.unit_embedding_(Wt.init_state_dict(Wt.iresnet100_spec(), seed=0))
    sr_sd = Wt.init_state_dict(Wt.sr_spec(), seed=0)
    d = make_inputs(0)
    n = C5_SRB
    a = (aei_sd, arc_sd, sr_sd, "linknet", 3, d["sources"][:n], d["crops"][:n], d["frames"][:n], d["masks"][:n],
         d["tfms"][:n])
    state = {}
    for _ in range(warmup):
        pipeline_cpu.swap_frames_sr_cpu(*a, sr_state=state)
    t0 = time.perf_counter()
    for _ in range(steps):
        pipeline_cpu.swap_frames_sr_cpu(*a, sr_state=state)
    dt = time.perf_counter() - t0
    sample = ("one SR batch of %d of the %d faces of a step per timed step (ArcFace + linknet/3 + LIPSPADEGenerator fp32 + "
              "cv2 paste-back into 1080p), %d timed step(s), %d warm-up, all host cores" % (n, C5_FACES, steps, warmup))
    return n * steps / dt, dt / steps, cores, sample


def run_reference_config5(args):
    steps, warm = max(1, args.steps), max(0, args.warmup)
    fps, sec, cores, sample = cpu_baseline_config5(steps, warm)
    emit({"impl": "reference", "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
          "warmup": warm, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
          "dtype": "f32", "data": "synthetic", "config": config5_config(args.gpus),
          "cpu_baseline": {"value": fps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
          "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0})
    return 0


def run_config5(args):
    """BASELINE.json configs[4]: AEI_Net('linknet', 3) bf16 -> pix2pix super-resolution generator
    (LIPSPADEGenerator, train-mode semantics, DataLoader batches of 20 as face_enhancement) -> paste-back of the
    enhanced crops into 1080p frames.  Reference-faithful variant: the SR generator is size-preserving (256 -> 256,
    SURVEY 8d config 5 (i)).  One step = 60 faces per GPU (three SR batches)."""
    import torch.distributed as dist
    import ghost_b200
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device (there is no CPU fallback)")
    cpu = cpu_baseline_config5(1, 0) if world == 1 else None   # alone on the host cores, before the GPU work
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # one process per GPU: sit on the CPUs / memory of the GPU's NUMA node before any pinned buffer is allocated
    from ghost_b200.shard import bind_to_gpu_numa
    numa = bind_to_gpu_numa(local) if os.environ.get("GHOST_NUMA_BIND", "1") != "0" else None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    pk = peaks()
    B5, SRB = C5_FACES, C5_SRB
    tdt = torch.float16 if args.dtype == "fp16" else torch.bfloat16
    G = ghost_b200.AEI_Net("linknet", 3, 512, seed=0).eval().to(tdt).cuda().enable_graph(True)
    arc = make_arcface(tdt).enable_graph(True)
    SR = ghost_b200.LIPSPADEGenerator().to(tdt).cuda().enable_graph(True)
    d_in = make_inputs(rank)
    sources_h, crops_h, frames_h, masks_h, tfms = (d_in[k] for k in ("sources", "crops", "frames", "masks", "tfms"))
    sl = slice(0, B5)
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    sources_p, crops_p, frames_p, masks_p = pin(sources_h[sl]), pin(crops_h[sl]), pin(frames_h[sl]), pin(masks_h[sl])
    out_p = torch.empty_like(frames_p).pin_memory()
    tf = tfms[sl]
    face_frame = np.arange(B5, dtype=np.int32)
    sources_d, crops_d, frames_d, masks_d = sources_p.to(dev), crops_p.to(dev), frames_p.to(dev), masks_p.to(dev)
    out_d = torch.empty_like(frames_d)
    swapped = torch.empty_like(crops_d)
    enhanced = torch.empty_like(crops_d)

    def step(src, crp, frm, msk, out):
        z = arc.embed_hwc(src, out_dtype=tdt)
        G.swap_u8(crp, z, out=swapped)
        for b0 in range(0, B5, SRB):
            SR.enhance_u8(swapped[b0:b0 + SRB], out=enhanced[b0:b0 + SRB])
        ghost_b200.paste_back(frm, enhanced, msk, tf, face_frame, mode="exact", out=out)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step(sources_d, crops_d, frames_d, masks_d, out_d)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step(sources_d, crops_d, frames_d, masks_d, out_d)
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    clocks = sampler.summary()
    # end to end: pinned host buffers in, composited frames back in the host buffer, every step - the streaming
    # pipeline with the SR stage, frames composited in place, only the face boxes over PCIe (as in the headline run)
    pipe = ghost_b200.SwapPipeline(G, arc, B5, (H, W), depth=2, mode="exact", boxes_only=True, sr=SR, sr_batch=SRB)
    def step_e2e():
        pipe.submit(sources_p, crops_p, frames_p, masks_p, tf, face_frame, frames_p)
    for _ in range(3):
        step_e2e()
    pipe.flush()
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    for _ in range(args.steps):
        step_e2e()
    pipe.flush()
    f1.record()
    barrier()
    ms_e2e = f0.elapsed_time(f1)
    t = torch.tensor([ms_total, ms_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_e2e = float(t[0]), float(t[1])
    line = None
    if rank == 0:
        prof = SR.profile(SRB, reps=3)
        conv = [o for o in prof if o["cls"] in (0, 1)]
        conv_ms, conv_fl = sum(o["ms"] for o in conv), sum(o["flops"] for o in conv)
        conv_tf = conv_fl / (conv_ms * 1e-3) / 1e12
        n_g, fl_g = G.stats(B5)
        n_a, fl_a = arc.stats(B5)
        n_s, fl_s = SR.stats(SRB)
        line = {
            "metric": METRIC, "value": B5 * world * args.steps / (ms_total * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f16" if tdt == torch.float16 else "bf16", "data": "synthetic",
            "config": config5_config(world), "host_numa_node": numa,
            "cpu_baseline": ({"value": cpu[0], "unit": UNIT, "cores": cpu[2], "kind": "port", "sample": cpu[3]}
                             if cpu is not None else
                             {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                              "sample": "measured at N=1 only (the other ranks would spin beside it)"}),
            "e2e": {"value": B5 * world * args.steps / (ms_e2e * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": int(pipe.h2d_bytes), "d2h_bytes_per_step": int(pipe.d2h_bytes),
                    "ms_per_step": ms_e2e / args.steps,
                    "note": "SwapPipeline(sr=SR, boxes_only=True): three streams, frames composited in place in the "
                            "pinned host buffer, only the face boxes cross PCIe"},
            "gpu_launches": int((n_g + n_a + 3 * n_s + 3 * 2 + 4) * args.steps),
            "clocks": clocks,
            "roofline": {"kernel": "igemm_kernel in the SR generator (tcgen05 convs incl. fused SPADE modulation, %d launches per batch of 20)" % len(conv),
                         "bound": "tensor", "achieved": conv_tf, "peak": pk["tf_sustained"], "unit": "TFLOP/s",
                         "frac": conv_tf / pk["tf_sustained"], "traffic": None,
                         "peak_source": pk["source"] + ", sustained",
                         "note": "algorithmic FLOPs of the unpadded layers (48/96/192-channel layers run padded to 64/128/256)"},
            "sr_ms_per_batch20": sum(o["ms"] for o in prof),
            "flops_per_step": fl_g + fl_a + 3 * fl_s,
        }
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        emit(line)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="default", choices=["default", "config4", "config5"],
                    help="default = BASELINE.json configs[1] + paste-back (the headline line); config4 = 4 faces per "
                         "frame (configs[3]); config5 = linknet/3 + SR (configs[4])")
    ap.add_argument("--dtype", default="fp16", choices=["fp16", "bf16"],
                    help="storage / operand type of the tcgen05 path: fp16 = the reference's G.half() (default), bf16")
    ap.add_argument("--no-eager", action="store_true", help="skip the reference_gpu_eager leg (PyTorch eager on the GPU)")
    args = ap.parse_args()
    quiet_stdout()
    if args.impl == "reference":
        return run_reference(args)
    if args.workload == "config5":
        return run_config5(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
