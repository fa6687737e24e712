#!/usr/bin/env python3
"""bench.py — throughput of the per-pixel Monte-Carlo loop (BASELINE.json metric: Mrays/s, samples/s) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--scene NAME] [--spp S]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

A STEP = one render of the workload scene at S samples per pixel PER GPU (bounce 5, sRGB on), i.e. one pass of the
hot path Renderer::ThreadRender -> PixelRender over the whole frame. With N GPUs every rank renders sample indices
[rank*S, (rank+1)*S) of an N*S-spp image into a full-frame fp32 {sum R, sum G, sum B, weight} buffer (weak scaling:
per-GPU work fixed), one NCCL reduce to rank 0 combines the buffers (the path's only exchange step, replacing the
MPI_Send/Recv gather of src/renderers/Renderer_MPI.cpp:142-207), and rank 0 resolves.
A RAY = one Scene::TraceNodeNormal or Scene::TraceNodeShadow call on the root node.

value   device-timed whole-job Mrays/s with the scene resident in HBM (CUDA events on the launching stream, max over
        ranks), accumulate + reduce only.
e2e     the same metric through the C ABI with HOST output buffers: qg_render (N=1) or accumulate/reduce/qg_resolve
        (N>1), frame planes copied device->host inside the timed region.
roofline  for the dominant kernel class (closest-hit or any-hit traversal): algorithmic bytes per the formula of
        SURVEY.md 8(d), counted by the -DQB_COUNT build of the same kernels, over the class's summed launch time.
cpu_baseline / --impl reference   the reference's own CPU renderer (oracle/_ref/qaref_perf = its sources compiled
        by oracle/Makefile, OpenMP tile loop) on this box's host cores, on a bounded spp sample of the same scene.
"""
import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

REF_INPUTS = os.path.join(ROOT, "oracle", "_ref", "inputs")
REF_PERF = os.path.join(ROOT, "oracle", "_ref", "qaref_perf")
ASSETS = os.path.join(ROOT, "assets")
METRIC = "Mrays/s"
BOUNCE = 5


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--scene", default="trc_scene_tower")  # the scene BASELINE.json's target is quoted on
    ap.add_argument("--spp", type=int, default=64, help="samples per pixel per GPU per step")
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--strong-spp", type=int, default=512, help="total spp of the strong-scaling sub-record (0 = skip); the same frame at every N")
    ap.add_argument("--tile-split", type=int, default=0, help="G_t of the strong job's G = G_t x G_s split (0 = 2 when N >= 4, else 1)")
    ap.add_argument("--sort-materials", action="store_true", help="qgRenderParams.sort_by_material = 1 (measured option, default off)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target length of the bounded CPU sample")
    return ap.parse_args()


def workload(args, width, height, world):
    """the `config` object: the same keys and values on the b200 arm and the reference arm (the driver compares them)"""
    return {"workload": "inputs/%s.xml (stand-in geometry from tools/gen_assets.py), %dx%d, %d spp/GPU/step, bounce %d, sRGB" % (
        args.scene, width, height, args.spp, BOUNCE), "scene": args.scene, "width": width, "height": height,
            "spp_per_gpu": args.spp, "spp_total": args.spp * world, "max_bounce": BOUNCE,
            "parallelism": "sample-shard x%d + NCCL reduce" % world if world > 1 else "1 GPU",
            "l2": "256 MiB device memset between steps (inside the timed region, ~0.1 ms); path-state working set also exceeds L2"}


def host_threads():
    """every host core this process may use. torch.distributed.run exports OMP_NUM_THREADS=1 to its workers: the reference arm
    must not inherit that (round-1 verdict: a 1-thread reference at N >= 2), so the count is passed to qaref_perf explicitly."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


# ------------------------------------------------------------------------------------------------- clocks
class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = str(gpu_index)
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", self.gpu], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
            self.thr.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if f[5 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------- CPU reference
def ref_available():
    return os.path.exists(REF_PERF) and os.path.exists(os.path.join(REF_INPUTS, "example_project3_sphere.xml"))


def ref_rays_per_sample(scene, spp=2):
    """closest + shadow rays per sample counted by the reference itself (oracle/_ref/libqaref_parity.so carries ray
    counters in Scene::TraceNode*, oracle/Makefile patch P3); stable to ~0.1 % across runs and spp."""
    code = ("import sys; sys.path.insert(0, %r)\nfrom oracle import refapi\nr = refapi.Ref(); r.set_params(%d, %d, %d, True, 0)\n"
            "r.load(%r, %r); r.counters(True); r.render(); c, s = r.counters(True); print('RAYS', c, s, r.W * r.H * %d)\n") % (
        ROOT, spp, spp, BOUNCE, os.path.join(REF_INPUTS, scene + ".xml"), ASSETS, spp)
    out = subprocess.run([sys.executable, "-c", code], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=900).stdout
    for line in out.splitlines():
        if line.startswith("RAYS"):
            _, c, s, n = line.split()
            return (int(c) + int(s)) / float(n), int(c) / float(n), int(s) / float(n)
    raise RuntimeError("reference ray counting failed: %s" % out[-500:])


def ref_perf_run(scene, spp, repeat, threads=None, dump_prefix=None):
    """the reference's own ThreadRender (its OpenMP tile loop, -O2 -march build) -> seconds per repeat, on ALL host cores;
    dump_prefix: repeat r's colour plane (rgb8) goes to <prefix><r>.rgb8"""
    threads = threads or host_threads()
    cmd = [REF_PERF, os.path.join(REF_INPUTS, scene + ".xml"), ASSETS, str(spp), str(BOUNCE), str(threads), str(repeat)]
    if dump_prefix:
        cmd.append(dump_prefix)
    env = dict(os.environ, OMP_NUM_THREADS=str(threads))
    env.pop("OMP_PROC_BIND", None)
    out = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=3600, env=env).stdout
    recs = [json.loads(l) for l in out.splitlines() if l.startswith("{")]
    if len(recs) != repeat:
        raise RuntimeError("qaref_perf failed: %s" % out[-500:])
    return recs


def cpu_baseline(scene, target_seconds, dump_prefix=None):
    """bounded sample: calibrate at 1 spp, then two back-to-back renders (the process's RNG streams run on, so the second is
    an independent estimate of the same frame) sized for about target_seconds together; both frames are kept for image_rmse"""
    rps, _, _ = ref_rays_per_sample(scene)
    cal = ref_perf_run(scene, 1, 1)[0]
    spp = int(max(1, min(256, round(0.5 * target_seconds / max(cal["seconds"], 1e-3)))))
    recs = ref_perf_run(scene, spp, 2, dump_prefix=dump_prefix)
    secs = sum(r["seconds"] for r in recs)
    samples = sum(r["samples"] for r in recs)
    mrays = samples * rps / secs * 1e-6
    rec = recs[0]
    return {"value": mrays, "unit": METRIC, "cores": rec["threads"], "host_cores": host_threads(), "kind": "reference",
            "sample": "%s at %dx%d, 2 x %d spp, bounce %d: %.2f s on %d OpenMP threads (reference's own ThreadRender, "
                      "oracle/_ref/qaref_perf); %.3f rays/sample counted by the reference" % (
                          scene, rec["width"], rec["height"], spp, BOUNCE, secs, rec["threads"], rps),
            "samples_per_s": samples / secs, "rays_per_sample": rps, "spp": spp, "width": rec["width"], "height": rec["height"]}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    if not ref_available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref not built (needs /root/reference at build time)"}))
        return 0
    rps, _, _ = ref_rays_per_sample(args.scene)
    cal = ref_perf_run(args.scene, 1, 1)[0]
    total = max(1, args.steps + args.warmup)
    spp = int(max(1, min(args.spp, round(150.0 / total / max(cal["seconds"], 1e-3)))))  # whole run within a few minutes
    recs = ref_perf_run(args.scene, spp, total)[args.warmup:]
    secs = sum(r["seconds"] for r in recs)
    samples = sum(r["samples"] for r in recs)
    val = samples * rps / secs * 1e-6
    cfg = workload(args, args.width or cal["width"], args.height or cal["height"], max(1, args.gpus))
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": METRIC, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": secs / len(recs) * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
            "samples_per_s": samples / secs,
            "reference_sample_spp": spp,
            "cpu_baseline": {"value": val, "unit": METRIC, "cores": recs[0]["threads"], "host_cores": host_threads(), "kind": "reference",
                             "sample": "%d spp per step (bounded sample of the %d spp workload), %.3f rays/sample" % (spp, args.spp, rps)},
            "e2e": {"value": val, "unit": METRIC, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------------- B200 arm
def algorithmic_bytes(closest, shadow, cst):
    """bytes(ray) = 32 (ray in) + 16 (hit out; 1 for any-hit) + 64 * BVH nodes visited + 48 * triangles
    tested + 48 * instance transforms applied (one per object test). cst = counting-build totals."""
    per_c = 32 + 16 + (64.0 * cst["nodes"] + 48.0 * cst["tris"] + 48.0 * cst["objects"]) / max(1, cst["closest_rays"])
    per_s = 32 + 1 + (64.0 * cst["s_nodes"] + 48.0 * cst["s_tris"] + 48.0 * cst["s_objects"]) / max(1, cst["shadow_rays"])
    return per_c * closest, per_s * shadow, per_c, per_s


def run_b200_arm(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from qaray_b200 import capi

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; qaray_b200 has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # NCCL_DEBUG stays as the operator set it (the rank count / transport lines are evidence); NCCL writes its log to
        # stdout by default, which carries the one JSON line, so the log goes to stderr unless a file was asked for
        if not os.environ.get("NCCL_DEBUG_FILE"):  # also without NCCL_DEBUG: the version banner goes where the log goes
            os.environ["NCCL_DEBUG_FILE"] = "/dev/stderr"
        dist.init_process_group("nccl", device_id=dev)
    lib = capi.load()
    capi.check(lib, lib.qg_init(local))
    comm = C.c_void_p()
    if world > 1:
        # the library's own communicator (qgComm over NCCL): torch.distributed only ships the 128-byte unique id and keeps the
        # harness's barriers / max-over-ranks; the timed shard render + reduce is qg_comm_render
        uid = torch.zeros(capi.QG_COMM_ID_BYTES, dtype=torch.uint8)
        if rank == 0:
            capi.check(lib, lib.qg_comm_unique_id(C.c_void_p(uid.data_ptr())))
        uid_d = uid.to(dev)
        dist.broadcast(uid_d, 0)
        uid = uid_d.cpu()
        capi.check(lib, lib.qg_comm_create(C.c_void_p(uid.data_ptr()), rank, world, C.byref(comm)))
    xml = os.path.join(REF_INPUTS, args.scene + ".xml")
    t_load = time.time()
    h = C.c_void_p()
    capi.check(lib, lib.qg_scene_load_xml(os.fsencode(xml), os.fsencode(ASSETS), C.byref(h)))
    t_load = time.time() - t_load
    info = capi.SceneInfo()
    capi.check(lib, lib.qg_scene_info(h, C.byref(info)))
    W = args.width or info.camera.width
    H = args.height or info.camera.height
    npix = W * H
    spp_total = args.spp * world

    def params(s0, s1, spp=spp_total):
        return capi.RenderParams(spp, BOUNCE, 1, 0x5EED, W, H, 0, 1, s0, s1, 0, 1 if args.sort_materials else 0)

    accum = torch.zeros(npix * 4, dtype=torch.float32, device=dev)
    depth = torch.zeros(npix, dtype=torch.float32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    # an explicit stream: a NULL stream handle means "the scene's own stream" to the library, which is not ordered with torch's
    # legacy default stream (the zero_() calls below)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    sptr = C.c_void_p(stream.cuda_stream)
    my = params(rank * args.spp, (rank + 1) * args.spp)
    whole = params(0, 0)  # the whole N*S-spp frame: qg_comm_render derives this rank's shard itself (sample split, tile_split 1)
    # host output planes (pinned) for the e2e path
    pin = lambda *shape, dt=torch.uint8: torch.empty(*shape, dtype=dt).pin_memory()
    h_mean, h_rgb8, h_depth = pin(npix * 3, dt=torch.float32), pin(npix * 3), pin(npix, dt=torch.float32)
    h_sc, h_mask = pin(npix), pin(npix)
    hp = lambda t: C.c_void_p(t.data_ptr())
    d2h_bytes = npix * (12 + 3 + 4 + 1 + 1)
    st = capi.RenderStats()
    tot = {"closest": 0, "shadow": 0, "samples": 0, "launches": 0, "trace_ms": 0.0, "shade_ms": 0.0, "shadow_ms": 0.0, "device_ms": 0.0}

    job = {"whole": whole, "tile_split": 1}  # what a step renders: the frame description all ranks share, and G_t

    def step(e2e, count):
        flush.zero_()  # L2 flush between steps
        my, whole, tile_split = job["whole"], job["whole"], job["tile_split"]
        if e2e and world == 1:
            capi.check(lib, lib.qg_render(h, C.byref(my), hp(h_mean), hp(h_rgb8), hp(h_depth), hp(h_sc), hp(h_mask), C.byref(st)))
        else:
            accum.zero_()
            depth.zero_()
            if world > 1:  # this rank's shard + ncclReduce(sum) of both planes to rank 0, inside the library
                capi.check(lib, lib.qg_comm_render(h, comm, C.byref(whole), tile_split, C.c_void_p(accum.data_ptr()), C.c_void_p(depth.data_ptr()), sptr,
                                                   C.byref(st)))
            else:
                capi.check(lib, lib.qg_render_accumulate(h, C.byref(my), C.c_void_p(accum.data_ptr()), C.c_void_p(depth.data_ptr()), sptr,
                                                         C.byref(st)))
            if e2e and rank == 0:
                capi.check(lib, lib.qg_resolve(h, C.byref(my), C.c_void_p(accum.data_ptr()), C.c_void_p(depth.data_ptr()), hp(h_mean),
                                               hp(h_rgb8), hp(h_depth), hp(h_sc), hp(h_mask), sptr))
        if count:
            tot["closest"] += st.closest_rays; tot["shadow"] += st.shadow_rays; tot["samples"] += st.samples
            tot["launches"] += st.kernel_launches + (1 if e2e and rank == 0 else 0)
            tot["trace_ms"] += st.trace_ms; tot["shade_ms"] += st.shade_ms; tot["shadow_ms"] += st.shadow_ms
            tot["device_ms"] += st.device_ms

    def timed(e2e, steps, sample_clocks):
        for k in tot:
            tot[k] = 0
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        clk = ClockSampler(local) if sample_clocks else None
        if clk:
            clk.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.time()
        e0.record(stream)
        for _ in range(steps):
            step(e2e, True)
        e1.record(stream)
        torch.cuda.synchronize()
        wall = time.time() - t0
        if world > 1:
            dist.barrier()
        clocks = clk.stop() if clk else None
        ms = e0.elapsed_time(e1)
        if e2e:
            ms = max(ms, wall * 1e3)  # host-side work (allocation, copies, waits) counts end to end
        v = torch.tensor([ms, float(tot["closest"] + tot["shadow"]), float(tot["samples"]), float(tot["launches"])], dtype=torch.float64, device=dev)
        if world > 1:
            mx = v.clone()
            dist.all_reduce(mx, op=dist.ReduceOp.MAX)
            dist.all_reduce(v, op=dist.ReduceOp.SUM)
            v[0] = mx[0]
        return {"ms": float(v[0]), "rays": float(v[1]), "samples": float(v[2]), "launches": int(v[3]), "clocks": clocks, "tot": dict(tot)}

    for _ in range(max(args.warmup, 3)):
        step(False, False)
    dev_run = timed(False, args.steps, True)
    step(True, False)
    e2e_run = timed(True, args.steps, False)
    # Strong scaling beside the weak line (the reference's MPI mode splits ONE frame over its ranks, renderer.cpp:383-398):
    # a fixed --strong-spp frame split tile x sample over the N ranks, G = G_t x G_s, device-timed like `value`.
    strong = None
    if args.strong_spp > 0 and args.strong_spp % world == 0:
        gt = args.tile_split if args.tile_split > 0 else (2 if world % 2 == 0 and world >= 4 else 1)
        job = {"whole": params(0, 0, spp=args.strong_spp), "tile_split": gt}
        step(False, False)
        sr = timed(False, max(1, min(args.steps, 3)), False)
        ns = max(1, min(args.steps, 3))
        strong = {"scaling": "strong", "spp_total": args.strong_spp, "split": "tiles x%d (interleaved 4x4 tiles) x samples x%d" % (gt, world // gt),
                  "steps": ns, "ms_per_step": sr["ms"] / ns, "value": sr["rays"] / (sr["ms"] * 1e-3) * 1e-6, "unit": METRIC,
                  "rays_per_step": sr["rays"] / ns, "n_gpus": world}
        job = {"whole": whole, "tile_split": 1}
        step(True, False)  # leave the weak job's reduced frame in accum for the image check below
    image_check = None
    if world > 1:
        # the reduced N x S-spp frame of the last step against the same frame rendered whole on rank 0's GPU (the RNG is keyed
        # by (pixel, sample), so only the order of the fp32 additions differs). Outside the timed region.
        red = accum.clone()
        if rank == 0:
            accum.zero_()
            depth.zero_()
            capi.check(lib, lib.qg_render_accumulate(h, C.byref(whole), C.c_void_p(accum.data_ptr()), C.c_void_p(depth.data_ptr()), sptr, None))
            torch.cuda.synchronize()
            a, b = red.view(-1, 4)[:, :3].double() / spp_total, accum.view(-1, 4)[:, :3].double() / spp_total
            # the reference's own arithmetic yields a few non-finite samples on the textured scenes (0/0 in its texture filter),
            # reproduced here: they must be the same pixels on both sides; the rest is compared
            fa, fb = torch.isfinite(a).all(dim=1), torch.isfinite(b).all(dim=1)
            m = fa & fb
            a, b = a[m], b[m]
            image_check = {"what": "NCCL-reduced %d-GPU frame vs the same %d-spp frame rendered on one GPU (linear radiance means)" % (world, spp_total),
                           "max_abs": float((a - b).abs().max()), "rmse": float(((a - b) ** 2).mean().sqrt()),
                           "max_rel": float(((a - b).abs() / (b.abs() + 1e-3)).max()), "mean": float(b.mean()),
                           "nonfinite_pixels": [int((~fa).sum()), int((~fb).sum())], "nonfinite_same_pixels": bool(torch.equal(fa, fb)),
                           "weights_equal": bool(torch.equal(red.view(-1, 4)[:, 3], accum.view(-1, 4)[:, 3]))}
        dist.barrier()
    return finish_b200(args, locals())


def image_rmse(lib, h, L, spp, prefix):
    """BASELINE.json `metric`: image RMSE. The GPU frame at the spp of the CPU baseline's bounded sample against the
    reference's frame of the same workload (the rgb8 plane PixelRender writes, renderer.cpp:348-361), beside the RMSE of
    the reference's two independent frames — the Monte-Carlo noise floor at that spp. Outside every timed region."""
    import numpy as np
    from qaray_b200 import capi
    W, H = L["W"], L["H"]
    a = np.fromfile(prefix + "0.rgb8", dtype=np.uint8).astype(np.float64) / 255.0
    b = np.fromfile(prefix + "1.rgb8", dtype=np.uint8).astype(np.float64) / 255.0
    g8 = np.zeros(W * H * 3, dtype=np.uint8)
    rp = capi.RenderParams(spp, BOUNCE, 1, 0x5EED, W, H, 0, 1, 0, 0, 0, 0)
    capi.check(lib, lib.qg_render(h, C.byref(rp), None, g8.ctypes.data_as(C.c_void_p), None, None, None, None))
    g = g8.astype(np.float64) / 255.0
    rm = lambda x, y: float(np.sqrt(np.mean((x - y) ** 2)))
    return {"gpu_vs_ref": rm(g, a), "gpu_vs_ref_b": rm(g, b), "ref_vs_ref": rm(a, b), "spp": spp, "width": W, "height": H,
            "space": "rgb8 plane / 255 (sRGB, clamped, as the reference saves it)",
            "mean_gpu": [float(v) for v in g.reshape(-1, 3).mean(axis=0)], "mean_ref": [float(v) for v in a.reshape(-1, 3).mean(axis=0)],
            "note": "independent Monte-Carlo estimates of one frame: gpu_vs_ref ~ ref_vs_ref means the GPU frame is as close "
                    "to the reference's as the reference is to itself"}


def finish_b200(args, L):
    """roofline (counting build), CPU baseline and the JSON line; rank 0 prints"""
    import torch
    import torch.distributed as dist
    from qaray_b200 import capi
    rank, world, lib, h = L["rank"], L["world"], L["lib"], L["h"]
    dev_run, e2e_run, info = L["dev_run"], L["e2e_run"], L["info"]
    out = None
    if rank == 0:
        # per-ray node / triangle / object-test counts from the -DQB_COUNT build of the same kernels, 4 spp of the same frame
        libc = capi.load(counting=True)
        capi.check(libc, libc.qg_init(L["local"]))
        hc = C.c_void_p()
        capi.check(libc, libc.qg_scene_load_xml(os.fsencode(L["xml"]), os.fsencode(ASSETS), C.byref(hc)))
        pc = capi.RenderParams(4, BOUNCE, 1, 0x5EED, L["W"], L["H"], 0, 1, 0, 0, 0, 0)
        sc = capi.RenderStats()
        acc = torch.zeros(L["npix"] * 4, dtype=torch.float32, device=L["dev"])
        capi.check(libc, libc.qg_render_accumulate(hc, C.byref(pc), C.c_void_p(acc.data_ptr()), None, L["sptr"], C.byref(sc)))
        libc.qg_scene_destroy(hc)
        cst = {"closest_rays": sc.closest_rays, "shadow_rays": sc.shadow_rays, "nodes": sc.nodes_visited, "tris": sc.tris_tested,
               "objects": sc.objects_tested, "s_nodes": sc.shadow_nodes_visited, "s_tris": sc.shadow_tris_tested,
               "s_objects": sc.shadow_objects_tested}
        t = dev_run["tot"]  # rank 0's own totals over the timed steps
        bytes_c, bytes_s, per_c, per_s = algorithmic_bytes(t["closest"], t["shadow"], cst)
        # The dominant kernel is the traversal core traverse_pool<ANY> in its two instantiations: k_wf_trace (closest hit) and
        # k_wf_shadow (any hit). shadow(d) runs beside trace(d + 1) on a second stream, so the two are timed as one: trace_ms +
        # shadow_ms is the CUDA-event time during which a traversal kernel was running (qgRenderStats), bytes are both kernels'.
        kname, kbytes, kms = "traverse_pool: k_wf_trace + k_wf_shadow (concurrent streams)", bytes_c + bytes_s, t["trace_ms"] + t["shadow_ms"]
        klaunch = kbytes / max(1.0, float(t["closest"] + t["shadow"]))
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
        else:
            peak, peak_src = 6650.0, "B200_PROFILING.md fallback (of fallback)"
        # launches per wave = 1 (stats) + (bounce + 1) x (trace + 2 cull-check kernels + shade [+ finish when the material sort
        # keeps it a separate pass] [+ shadow + its exact walk]); waves per step follow
        # (finish runs at depth 0 only: deeper depths use the fused shade kernel; with the material sort every depth has it, plus 3 sort kernels)
        per_wave = 1 + (BOUNCE + 1) * (4 + (2 if t["shadow"] > 0 else 0)) + (4 * (BOUNCE + 1) if args.sort_materials else 1)
        n_waves = max(1, int(round(t["launches"] / float(args.steps) / per_wave)))
        achieved = kbytes / (kms * 1e-3) * 1e-9 if kms > 0 else 0.0
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")  # DRAM bytes per ray of the kernel from the ncu --set full capture
        klaunches = max(1, (BOUNCE + 1) * args.steps * n_waves * (2 if t["shadow"] > 0 else 1))
        if os.path.exists(tp):
            rec = json.load(open(tp))
            if "k_wf_trace" in rec and "k_wf_shadow" in rec:  # per launch, like achieved's bytes
                traffic = (rec["k_wf_trace"]["dram_bytes_per_ray"] * t["closest"] + rec["k_wf_shadow"]["dram_bytes_per_ray"] * t["shadow"]) / klaunches
        roofline = {"bound": "hbm", "kernel": kname, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                    "traffic": traffic, "peak_source": peak_src, "bytes_per_ray": klaunch, "kernel_ms_per_step": kms / args.steps,
                    "launches_of_kernel": klaunches, "algorithmic_bytes_per_launch": kbytes / klaunches,
                    "per_ray": {"closest": {"nodes": cst["nodes"] / max(1, cst["closest_rays"]), "tris": cst["tris"] / max(1, cst["closest_rays"]),
                                            "objects": cst["objects"] / max(1, cst["closest_rays"]), "bytes": per_c},
                                "shadow": {"nodes": cst["s_nodes"] / max(1, cst["shadow_rays"]), "tris": cst["s_tris"] / max(1, cst["shadow_rays"]),
                                           "objects": cst["s_objects"] / max(1, cst["shadow_rays"]), "bytes": per_s}},
                    "class_ms_per_step": {"trace": t["trace_ms"] / args.steps, "shade": t["shade_ms"] / args.steps,
                                          "shadow": t["shadow_ms"] / args.steps, "render_total": t["device_ms"] / args.steps}}
        value = dev_run["rays"] / (dev_run["ms"] * 1e-3) * 1e-6
        e2e_v = e2e_run["rays"] / (e2e_run["ms"] * 1e-3) * 1e-6
        cfg = workload(args, L["W"], L["H"], world)
        scene_stats = {"triangles": int(info.num_faces), "bvh_nodes": int(info.num_bvh_nodes), "scene_device_bytes": int(info.device_bytes),
                       "scene_load_s": L["t_load"], "rays_per_sample": dev_run["rays"] / max(1.0, dev_run["samples"])}
        out = {"metric": METRIC, "value": value, "unit": METRIC, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
               "ms_per_step": dev_run["ms"] / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
               "dtype": "f32", "data": "synthetic", "config": cfg,
               "samples_per_s": dev_run["samples"] / (dev_run["ms"] * 1e-3),
               "samples_per_s_per_gpu": dev_run["samples"] / (dev_run["ms"] * 1e-3) / world,
               "clocks": dev_run["clocks"], "gpu_launches": dev_run["launches"],
               "e2e": {"value": e2e_v, "unit": METRIC, "h2d_bytes_per_step": C.sizeof(capi.RenderParams),
                       "d2h_bytes_per_step": L["d2h_bytes"], "ms_per_step": e2e_run["ms"] / args.steps,
                       "note": "scene resident (uploaded once by qg_scene_load_xml, as the reference loads its scene before Render); "
                               "per-step input is the parameter block; output = rgb_mean f32 + rgb8 + depth f32 + sampleCount + mask"},
               "roofline": roofline, "scene_stats": scene_stats}
        if L.get("strong") is not None:
            out["strong"] = L["strong"]
        if L.get("image_check") is not None:
            out["image_check"] = L["image_check"]
        if world == 1 and not args.no_cpu_baseline and ref_available():
            try:
                import tempfile
                with tempfile.TemporaryDirectory() as td:
                    prefix = os.path.join(td, "ref_")
                    out["cpu_baseline"] = cb = cpu_baseline(args.scene, args.cpu_seconds, prefix)
                    if cb["width"] == L["W"] and cb["height"] == L["H"]:
                        out["image_rmse"] = image_rmse(lib, h, L, cb["spp"], prefix)
            except Exception as e:  # the baseline is reported, never required for the GPU number
                out["cpu_baseline"] = {"value": None, "unit": METRIC, "cores": 0, "kind": "reference", "sample": "failed: %s" % e}
        print(json.dumps(out))
    lib.qg_scene_destroy(h)
    if world > 1:
        lib.qg_comm_destroy(L["comm"])
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    args = parse_args()
    # stdout carries the ONE JSON line and nothing else: libraries that print there (NCCL's version banner ignores
    # NCCL_DEBUG_FILE) are pointed at stderr at the descriptor level, and Python's stdout keeps the original descriptor
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w", buffering=1)
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_b200_arm(args)


if __name__ == "__main__":
    sys.exit(main())
