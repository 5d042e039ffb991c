#!/usr/bin/env python
"""bench.py -- whole-frame JPEG 2000 decode throughput of the B200 tile decoder.

Workload (BASELINE.json configs[1]): 4096x2160 16-bit RGB, irreversible 9/7 + ICT, 5 decomposition
levels, 64x64 code blocks, one tile, one layer, every coding pass kept -- the codestream the STOCK
reference encoder writes for the synthetic frame.  A step = one decode of that codestream:

    ours       value = device-resident: T1 (EBCOT/MQ + fused dequantisation) + inverse 9/7 of three
                       components + fused inverse ICT / DC shift / clamp, tables already in HBM,
                       CUDA events on the decoder's stream;
               e2e   = the C-ABI codestream door (grkb_decompress_submit / _wait): coded bytes in
                       page-locked host memory in, our own marker + packet parsing (T2) on the host,
                       H2D, all kernels, D2H of the int32 planes -- the same work the reference arm
                       does on its memory stream; K frames in flight;
               e2e_plugin / e2e_plugin_batch (N = 1) = the same frame through the reference's own
                       plugin doors, grk_plugin_decompress and grk_plugin_batch_decompress, driven by
                       the UNMODIFIED reference host (oracle/_ref) that loads libgrokj2k_plugin.so;
               ht    = the same three figures for BASELINE configs[2] (12-bit HTJ2K lossless);
               parity = GPU planes compared with the stock decoder's planes of the same codestream.
    reference  the unmodified reference's own multithreaded CPU decoder (public C API, memory
               stream, oracle/_ref/libgrok_ref.so) on the same codestream, all host threads.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload j2k97|ht]

Synthetic data; frames are sharded over ranks with no collective (weak scaling).
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

W, H, NCOMP, NUMRES, CBLK = 4096, 2160, 3, 6, 6
SEED = 4096
WORKLOADS = {
    "j2k97": dict(name="j2k_decode_4096x2160_rgb16_97_ict_5lvl_cblk64_1tile", prec=16, ht=False, irr=1,
                  dtype="f32"),
    "ht": dict(name="htj2k_decode_4096x2160_rgb12_53_rct_lossless_5lvl_cblk64_1tile", prec=12, ht=True, irr=0,
               dtype="i32"),
}
# MQ decisions of the configs[1] frame and the kernel's instruction counts: from the committed ncu
# capture, not measured in this run
T1_PROFILE = os.path.join(ROOT, "profiles", "t1_mq_decode_profile.json")


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def dist_env():
    """(rank, local_rank, world_size) from the torchrun environment"""
    rank = int(os.environ.get("RANK", "0"))
    return rank, int(os.environ.get("LOCAL_RANK", str(rank))), int(os.environ.get("WORLD_SIZE", "1"))


def config_dict(wl, world):
    """identical for both arms (anything else goes under "detail")"""
    return {"workload": wl["name"], "width": W, "height": H, "components": NCOMP, "precision": wl["prec"],
            "wavelet": "5/3" if wl["ht"] else "9/7", "mct": "RCT" if wl["ht"] else "ICT", "levels": NUMRES - 1,
            "block_coder": "HT" if wl["ht"] else "MQ", "cblk": "64x64", "frames_per_step": world}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def cache_path(tag, wl):
    key = hashlib.sha1(("%s-%d-%d-%d-%d-%d-%d-v5" % (tag, W, H, wl["prec"], NUMRES, SEED, wl["ht"])).encode()).hexdigest()[:12]
    return os.path.join(os.environ.get("GRKB_BENCH_CACHE", "/tmp"), "grkb_bench_%s_%s.npz" % (tag, key))


def synth_frame(wl):
    """the 16-bit frame of configs[1]; configs[2] is the same field >> 4"""
    import synth as Hh
    base = dict(WORKLOADS["j2k97"])
    p = cache_path("img", base)
    if os.path.exists(p):
        img = list(np.load(p)["img"])
    else:
        t = time.time()
        img = Hh.synth_image(W, H, NCOMP, 16, SEED)
        np.savez(p + ".tmp.npz", img=np.stack(img))
        os.replace(p + ".tmp.npz", p)
        log("[bench] synthetic frame generated in %.1f s" % (time.time() - t))
    return [a >> (16 - wl["prec"]) for a in img] if wl["prec"] != 16 else img


def codestream(wl, img=None):
    """the stock reference encoder's codestream of the frame (set-up, untimed)"""
    from oracle import oracle as O
    p = cache_path("cs", wl)
    if os.path.exists(p):
        return np.load(p)["cs"]
    if not O.have_ref():
        raise SystemExit("bench.py: oracle/_ref/libgrok_ref.so (the compiled reference) is needed to write the "
                         "workload's codestream; build it with __graft_entry__.build() where /root/reference exists")
    if img is None:
        img = synth_frame(wl)
    t = time.time()
    cs = O.ref_codec_encode(img, wl["prec"], wl["irr"], ht=wl["ht"], numres=NUMRES, cblk=(1 << CBLK, 1 << CBLK),
                            threads=os.cpu_count())
    np.savez(p + ".tmp.npz", cs=cs)
    os.replace(p + ".tmp.npz", p)
    log("[bench] reference codestream written in %.1f s: %.1f MB" % (time.time() - t, cs.size / 1e6))
    return cs


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region"""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4)
                          if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------------------ reference arm

def cpu_reference_decode(cs, steps, warmup, threads):
    """the stock reference decoder on the whole frame; returns (MP/s, seconds per frame)"""
    from oracle import oracle as O
    for _ in range(warmup):
        O.ref_codec_decode(cs, NCOMP, W, H, threads=threads, want_planes=False)
    t = time.perf_counter()
    for _ in range(steps):
        O.ref_codec_decode(cs, NCOMP, W, H, threads=threads, want_planes=False)
    dt = (time.perf_counter() - t) / steps
    return W * H / 1e6 / dt, dt


def cli_timings(wl):
    """the stock grk_decompress BINARY (oracle/_ref/bin, built from the
    reference's own CLI sources) on the workload's codestream, wall clock of the whole process
    (file read, decode, 16-bit PPM write), all host threads and one"""
    exe = os.path.join(ROOT, "oracle", "_ref", "bin", "grk_decompress")
    if not os.path.exists(exe):
        return None
    import tempfile
    d = tempfile.mkdtemp(prefix="grkb_cli_")
    try:
        src, dst = os.path.join(d, "f.j2k"), os.path.join(d, "f.ppm")
        codestream(wl).tofile(src)
        out = {}
        for label, h in (("H_all", os.cpu_count() or 1), ("H_1", 1)):
            t = time.perf_counter()
            r = subprocess.run([exe, "-i", src, "-o", dst, "-H", str(h)], stdout=subprocess.DEVNULL,
                               stderr=subprocess.DEVNULL, timeout=300)
            if r.returncode != 0:
                return {"error": "grk_decompress exited %d" % r.returncode}
            out["grk_decompress_%s_ms" % label] = (time.perf_counter() - t) * 1e3
        out["what"] = "stock CLI binary, whole process wall clock incl. file read and PPM write; H_all = %d threads" % (
            os.cpu_count() or 1)
        return out
    except Exception as e:
        return {"error": str(e)[:120]}
    finally:
        import shutil
        shutil.rmtree(d, ignore_errors=True)


def run_reference(args, wl):
    rank, local, world = dist_env()
    if rank != 0:
        return 0
    threads = os.cpu_count() or 1
    cs = codestream(wl)
    val, dt = cpu_reference_decode(cs, args.steps, args.warmup, threads)
    line = {"impl": "reference", "metric": "decode_mpixels_per_s", "value": val, "unit": "Mpixel/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": wl["dtype"],
            "data": "synthetic", "gpu_launches": 0, "config": config_dict(wl, world),
            "cpu_baseline": {"value": val, "unit": "Mpixel/s", "cores": threads, "kind": "reference",
                             "sample": "whole frame: stock grk decompress (public C API, memory stream, markers + T2 "
                                       "+ T1 + wavelet + MCT), %d steps" % args.steps},
            "e2e": {"value": val, "unit": "Mpixel/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------ our arm

def measure(wl, args, local, barrier, with_parity=True):
    """all legs of one workload on this rank's GPU; returns a dict of raw figures (ms per frame)"""
    from grok_b200 import codestream as CS
    from grok_b200 import tile as T
    from oracle import oracle as O
    cs_arr = codestream(wl)
    pin_cs = T.PinnedArray((cs_arr.size + 16,), np.uint8)
    pin_cs.array[:cs_arr.size] = cs_arr
    cs = pin_cs.array[:cs_arr.size]
    # ---- our own T2 once, for the device-resident leg
    c = CS.Codestream(cs)
    t0 = time.perf_counter()
    tab = c.parse_tile(0)
    t2_ms = (time.perf_counter() - t0) * 1e3
    c.close()
    blocks, segs, data = tab["blocks"], tab["segs"], tab["data"]
    params = tab["params"]
    dec = T.TileDecoder(params, len(blocks), data.size)
    pin_out = [T.PinnedArray((H, W), np.int32) for _ in range(NCOMP)]
    out = [p.array for p in pin_out]
    dec.decode(blocks, segs, data, out=out)
    if dec.rejected:
        raise SystemExit("bench.py: %d code blocks rejected" % dec.rejected)
    res = {"blocks": int(len(blocks)), "coded_bytes": int(cs_arr.size), "t2_ms_single": t2_ms,
           "passes_max": int(segs["numpasses"].max()) if len(segs) else 0}
    # ---- parity: the stock decoder's planes of the very same codestream (untimed)
    if with_parity:
        if O.have_ref():
            want = O.ref_codec_decode(cs_arr, NCOMP, W, H, threads=os.cpu_count())
            diff = max(int(np.abs(a.astype(np.int64) - b).max()) for a, b in zip(out, want))
            res["parity"] = {"checked": True, "max_abs_diff": diff, "vs": "oracle/_ref stock grk_decompress of the same "
                             "codestream", "tolerance": 0 if wl["ht"] else 1}
            if diff > (0 if wl["ht"] else 1):
                raise SystemExit("bench.py: PARITY FAILURE on %s: max |diff| = %d" % (wl["name"], diff))
        else:
            res["parity"] = {"checked": False, "why": "oracle/_ref not present on this box"}
    # ---- value: device-resident decode, CUDA events on the decoder's stream
    barrier()
    res["tm"] = dec.time_device(args.steps, args.warmup)
    barrier()
    # ---- the same with four frames in flight (a decoder and a stream each, everything resident): what the
    # device sustains when a frame's latency is not the point -- the block decoder then pairs code blocks
    # (an extra, N = 1 only: no collective inside, and a failure costs the key, not the run)
    if not wl["ht"] and dist_env()[2] == 1:
        more = []
        try:
            import threading
            more = [T.TileDecoder(params, len(blocks), data.size) for _ in range(3)]
            for m in more:
                m.decode(blocks, segs, data, out=out)
            group = [dec] + more
            n_each = max(3, args.steps // 2)

            def spin(d_):
                d_.time_device(n_each, 1)
            ths = [threading.Thread(target=spin, args=(d_,)) for d_ in group]
            t0 = time.perf_counter()
            for t_ in ths:
                t_.start()
            for t_ in ths:
                t_.join()
            # (grkb_decoder_time_device runs its warm-up, the timed passes, and the same passes again with events
            # around the stages: 2 x n_each + 1 frames per decoder)
            res["inflight4_ms"] = (time.perf_counter() - t0) * 1e3 / (len(group) * (2 * n_each + 1))
        except Exception as e:
            log("[bench] value_4_frames_in_flight skipped: %s" % str(e)[:200])
        for m in more:
            try:
                m.close()
            except Exception:
                pass
    barrier()
    # ---- e2e: the codestream door, K frames in flight (one decompressor each, one slot each)
    K = max(1, args.inflight if args.inflight else (2 if wl["ht"] else 4))
    dzs = [CS.Decompressor(local, slots=1) for _ in range(K)]
    outs = [out] + [[T.PinnedArray((H, W), np.int32).array for _ in range(NCOMP)] for _ in range(K - 1)]

    def e2e_pass(n, packed=None):
        for i in range(n):
            d = dzs[i % K]
            if i >= K:
                d.wait()
            if packed is None:
                d.submit(cs, outs[i % K])
            else:
                d.submit_packed(cs, packed[i % K], NCOMP)
        for i in range(max(n - K, 0), n):
            dzs[i % K].wait()

    for o in outs:
        for p in o:
            p[:] = 0
    e2e_pass(max(min(args.warmup, 3), K))
    barrier()
    t0 = time.perf_counter()
    e2e_pass(args.steps)
    res["e2e_ms"] = (time.perf_counter() - t0) * 1e3 / args.steps
    barrier()
    res["t2_ms"] = dzs[0].stats()["t2_ms"]
    if any(d.rejected for d in dzs) or not all(np.array_equal(a, b) for o in outs[1:] for a, b in zip(outs[0], o)):
        raise SystemExit("bench.py: the frames in flight disagree")
    # ---- the same with 16-bit samples across the bus (row N3, opt-in)
    outs16 = [T.PinnedArray((NCOMP, H, W), np.uint16).array for _ in range(K)]
    e2e_pass(max(min(args.warmup, 3), K), outs16)
    barrier()
    t0 = time.perf_counter()
    e2e_pass(args.steps, outs16)
    res["e2e16_ms"] = (time.perf_counter() - t0) * 1e3 / args.steps
    barrier()
    if not all(np.array_equal(o[k], outs[0][k].astype(np.uint16)) for o in outs16 for k in range(NCOMP)):
        raise SystemExit("bench.py: 16-bit output differs from the int32 planes")
    res["K"] = K
    res["h2d"] = int(cs_arr.size + blocks.nbytes + segs.nbytes + 4 * len(blocks))
    for d in dzs:
        d.close()
    dec.close()
    return res


_PLUGIN_OPEN = False


def plugin_legs(wl, args):
    """N = 1: the frame through the reference's own plugin doors with the unmodified host"""
    from oracle import oracle as O
    import tempfile
    if not O.have_plugin():
        return None
    try:
        import ctypes as C
        global _PLUGIN_OPEN
        if not _PLUGIN_OPEN:
            O.ref_plugin_open(device=0)
            _PLUGIN_OPEN = True
        cs_arr = codestream(wl)
        d = tempfile.mkdtemp(prefix="grkb_bench_")
        path = os.path.join(d, "frame.j2k")
        cs_arr.tofile(path)
        ms = C.c_double()
        steps = max(3, min(args.steps, 10))
        rc = O.ref().ref_plugin_decode_bench(path.encode(), C.c_uint32(NCOMP), C.c_uint32(W), C.c_uint32(H),
                                            C.c_uint32(steps), C.c_uint32(2), C.byref(ms))
        out = {}
        if rc == 0:
            out["e2e_plugin"] = {"value": W * H / 1e6 / (ms.value * 1e-3), "unit": "Mpixel/s", "ms_per_step": ms.value,
                                 "door": "grk_plugin_decompress, one synchronous call per frame: file read + host "
                                         "header + own T2 + H2D + kernels + D2H into the host's grk_image",
                                 "host": "unmodified reference (oracle/_ref) with libgrokj2k_plugin.so loaded"}
        else:
            out["e2e_plugin"] = {"value": None, "error": "ref_plugin_decode_bench rc %d" % rc}
        nb = 16
        ind, outd = os.path.join(d, "in"), os.path.join(d, "out")
        os.makedirs(ind)
        os.makedirs(outd)
        for i in range(nb):
            os.link(path, os.path.join(ind, "f%03d.j2k" % i))
        O.ref().ref_plugin_batch_decode_bench(ind.encode(), outd.encode(), C.byref(ms))  # warm-up
        n = O.ref().ref_plugin_batch_decode_bench(ind.encode(), outd.encode(), C.byref(ms))
        if n == nb:
            out["e2e_plugin_batch"] = {"value": nb * W * H / 1e6 / (ms.value * 1e-3), "unit": "Mpixel/s",
                                       "ms_per_step": ms.value / nb, "frames": nb,
                                       "door": "grk_plugin_batch_decompress over a directory of %d files, the plugin's "
                                               "worker pool (one decoder per worker)" % nb}
        else:
            out["e2e_plugin_batch"] = {"value": None, "error": "batch returned %d" % n}
        import shutil
        shutil.rmtree(d, ignore_errors=True)
        return out
    except Exception as e:  # the plugin legs are extra keys; the line stands without them
        return {"e2e_plugin": {"value": None, "error": str(e)[:200]}}


ENCODE_LEGS = ("dci4k_97_mq_cblk32_rate_cinema4k", "dci4k_97_mq_cblk32_rate_cinema4k_cprl_prc256", "ht_lossless_cblk64")
T8K, TILE8K = 8192, 1024


def frame8k():
    """BASELINE configs[3] input: 8192x8192x3 16-bit, built from one 2048x2048 synthetic field repeated
    with a different shift per 2048-block (a full-size synth costs 25 s)"""
    import synth as Hh
    base = Hh.synth_image(2048, 2048, NCOMP, 16, 8192)
    out = []
    for c, p in enumerate(base):
        rows = [np.concatenate([np.roll(p, (97 * (i + 4 * j + c), 31 * (i + j)), axis=(0, 1)) ^ (i * 257 + j)
                                for i in range(4)], axis=1) for j in range(4)]
        out.append(np.ascontiguousarray(np.concatenate(rows, axis=0)).astype(np.int32))
    return out


def codestream8k():
    from oracle import oracle as O
    wl = dict(prec=16, ht=True)
    p = cache_path("cs8k", wl)
    if os.path.exists(p):
        return np.load(p)["cs"]
    if not O.have_ref():
        return None
    t = time.time()
    cs = O.ref_codec_encode(frame8k(), 16, 0, ht=True, numres=NUMRES, cblk=(64, 64), tile=(TILE8K, TILE8K),
                            threads=os.cpu_count())
    np.savez(p + ".tmp.npz", cs=cs)
    os.replace(p + ".tmp.npz", p)
    log("[bench] 8192x8192 tiled HTJ2K codestream written in %.1f s: %.1f MB" % (time.time() - t, cs.size / 1e6))
    return cs


def tiled_leg(args, local, rank, world, barrier):
    """BASELINE configs[3]: 8192x8192 16-bit RGB HTJ2K in 64 tiles of 1024x1024, tile t on rank t mod N
    (grkb_decompress_submit_tiles, no collective); codestream bytes in page-locked host memory in, this
    rank's tiles as int32 rectangles of a page-locked image out, own T2 inside.  Returns this rank's
    milliseconds per image (the ranks work on the same image: strong scaling)."""
    from grok_b200 import codestream as CS
    from grok_b200 import tile as T
    cs_arr = codestream8k()
    if cs_arr is None:
        return None
    pin = T.PinnedArray((cs_arr.size + 16,), np.uint8)
    pin.array[:cs_arr.size] = cs_arr
    cs = pin.array[:cs_arr.size]
    out = [T.PinnedArray((T8K, T8K), np.int32).array for _ in range(NCOMP)]
    dz = CS.Decompressor(local, slots=4)
    steps = max(3, args.steps // 4)
    for _ in range(2):
        dz.submit_tiles(cs, out, rank, world)
        dz.wait()
    t0 = time.perf_counter()
    for _ in range(steps):
        dz.submit_tiles(cs, out, rank, world)
        dz.wait()
    ms = (time.perf_counter() - t0) * 1e3 / steps
    res = {"ms": ms, "rejected": dz.rejected, "tiles_this_rank": dz.stats()["tiles"], "coded_bytes": int(cs_arr.size)}
    if rank == 0:
        # lossless: this rank's tiles must be the image
        img = frame8k()
        nt = T8K // TILE8K
        ok = True
        for t in range(rank, nt * nt, world):
            y, x = (t // nt) * TILE8K, (t % nt) * TILE8K
            ok = ok and all(np.array_equal(out[c][y:y + TILE8K, x:x + TILE8K], img[c][y:y + TILE8K, x:x + TILE8K])
                            for c in range(NCOMP))
        res["parity_ok"] = bool(ok)
        if not ok:
            raise SystemExit("bench.py: PARITY FAILURE on the tiled 8192x8192 leg")
    dz.close()
    return res


def encode_legs(args, local=0, world=1):
    """N = 1 extras: the ENCODE side of the codestream door (grkb_compress) on the frames of BASELINE
    configs[4] (DCI-4K 12-bit 9/7, 32x32 blocks, one layer at the cinema-4K byte budget: T1 with
    distortions on the device, the reference's PCRD bisection + T2 + markers on the host) and
    configs[2] (HTJ2K lossless), several host threads with a compressor each; wall clock."""
    try:
        from grok_b200 import codestream as CS
        from grok_b200 import tile as T
        from oracle import oracle as O
        img = [a >> 4 for a in synth_frame(WORKLOADS["j2k97"])]
        pins = [T.PinnedArray((H, W), np.int32) for _ in range(NCOMP)]
        for p_, b in zip(pins, img):
            p_.array[:] = b
        planes = [p_.array for p_ in pins]
        out = {}
        nthreads = max(int(os.environ.get("GRKB_BENCH_ENCODE_THREADS", "2")), min(16, (os.cpu_count() or 2) // world))
        cinema_rate = W * H * 3 * 1.5 / 1302083.0  # the cinema-4K byte budget per frame at 24 fps
        cinema_prc = [(256, 256)] * (NUMRES - 1) + [(128, 128)]  # and the profile's precincts and progression
        for name, kw, nthr in (("dci4k_97_mq_cblk32_rate_cinema4k", dict(reversible=False, cblk=(32, 32),
                                                                       rate=cinema_rate), nthreads),
                               ("dci4k_97_mq_cblk32_rate_cinema4k_cprl_prc256",
                                dict(reversible=False, cblk=(32, 32), rate=cinema_rate, prog="CPRL", precincts=cinema_prc),
                                nthreads),
                               ("ht_lossless_cblk64", dict(reversible=True, ht=True, cblk=(64, 64)), min(8, nthreads))):
            zs = [CS.Compressor(W, H, NCOMP, 12, numres=NUMRES, comment="Created by Grok     version 9.2.0", device=local,
                                **kw) for _ in range(nthr)]
            first = [z.encode(planes) for z in zs]
            per = 2
            ths = [threading.Thread(target=lambda z=z: [z.encode(planes) for _ in range(per)]) for z in zs]
            t0 = time.perf_counter()
            for t in ths:
                t.start()
            for t in ths:
                t.join()
            dt = (time.perf_counter() - t0) / (per * nthr)
            st = zs[0].stats()
            for z in zs:
                z.close()
            leg = {"value": W * H / 1e6 / dt, "unit": "Mpixel/s", "ms_per_frame": dt * 1e3, "host_threads": nthr,
                   "coded_bytes": int(first[0].size),
                   "one_frame_alone_ms": {"device_incl_copies": st["device_ms"], "host_t2_and_rate_allocation": st["t2_ms"]}}
            if O.have_ref() and local == 0:
                try:
                    O.ref_set_rate(kw.get("rate", 0))
                    if "precincts" in kw:
                        O.ref_set_encode_options("prog=%s;prc=%s" % (kw["prog"], ":".join("%dx%d" % p_ for p_ in kw["precincts"])))
                    t0 = time.perf_counter()
                    ref = O.ref_codec_encode(img, 12, 0 if kw.get("ht") else 1, ht=bool(kw.get("ht")), numres=NUMRES,
                                             cblk=kw["cblk"], threads=os.cpu_count())
                    ct = time.perf_counter() - t0
                finally:
                    O.ref_set_rate(0)
                    O.ref_set_encode_options("")
                leg["parity"] = {"byte_identical_to_stock_encoder": bool(ref.size == first[0].size and
                                                                         np.array_equal(ref, first[0]))}
                leg["cpu_baseline"] = {"value": W * H / 1e6 / ct, "unit": "Mpixel/s", "cores": os.cpu_count(),
                                       "kind": "reference", "sample": "one frame, stock encoder, memory stream, %.0f ms" % (ct * 1e3)}
                if not leg["parity"]["byte_identical_to_stock_encoder"]:
                    raise SystemExit("bench.py: PARITY FAILURE: encode leg %s differs from the stock encoder" % name)
            out[name] = leg
        return out
    except SystemExit:
        raise
    except Exception as e:  # extra keys only
        return {"error": str(e)[:200]}


_REAL_STDOUT = None
COLLECTIVE_BACKEND, COLLECTIVE_DEVICE = "nccl", "cuda"  # (tests/test_bench_flow.py runs the N > 1 flow over gloo)


def emit(line):
    """the one JSON line, on the process's real standard output"""
    sys.stdout.flush()
    if _REAL_STDOUT is not None:
        os.dup2(_REAL_STDOUT, 1)
    print(json.dumps(line), flush=True)


def run_ours(args, wl):
    rank, local, world = dist_env()
    import grok_b200
    from grok_b200 import shard
    from oracle import oracle as O
    if grok_b200.lib.grkb_device_count() < 1:
        raise SystemExit("bench.py: no CUDA device visible; the product has no CPU fallback")
    grok_b200._lib.check(grok_b200.lib.grkb_set_device(local), "grkb_set_device")
    numa_bound = shard.bind_to_device(local) if world > 1 else False
    dist = None
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"  # keeps NCCL's banner off stdout: one JSON line only
        # ... and whatever a library still writes to file descriptor 1 goes to stderr until the line is printed
        global _REAL_STDOUT
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)
        import torch
        import torch.distributed as dist
        import datetime
        # (a rank that dies takes the others down through torchrun; a rank that waits for one that never comes
        # gives up after five minutes instead of NCCL's ten)
        if COLLECTIVE_BACKEND == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group("nccl", device_id=torch.device("cuda", local),
                                    timeout=datetime.timedelta(seconds=300))
        else:
            dist.init_process_group(COLLECTIVE_BACKEND, timeout=datetime.timedelta(seconds=300))

    def barrier():
        if dist is not None:
            dist.barrier()

    # ---- set-up (untimed): rank 0 writes the caches, the others read them
    wl_ht = WORKLOADS["ht"]
    second = None if (wl["ht"] or args.no_ht) else wl_ht
    if rank == 0:
        codestream(wl)
        if second:
            codestream(second)
        if not args.no_tiled:
            codestream8k()
    barrier()
    sampler = ClockSampler(local)
    main = measure(wl, args, local, barrier)
    sub = None
    if second:
        a2 = argparse.Namespace(**vars(args))
        a2.steps = max(5, args.steps // 2)
        sub = measure(second, a2, local, barrier)
    # ---- extras on every rank: configs[3] tile-sharded over the ranks, configs[4] / [2] encode frame-sharded
    # (no collective inside them; a leg that fails on a rank counts as zero there and is reported)
    tiled = None
    if not args.no_tiled:
        try:
            tiled = tiled_leg(args, local, rank, world, barrier)
        except SystemExit:
            raise
        except Exception as e:
            log("[bench] rank %d: tiled_8k failed: %s" % (rank, str(e)[:200]))
    barrier()
    enc = None if args.no_encode else encode_legs(args, local, world)
    barrier()
    # ---- the link, all ranks at once
    import ctypes as C
    barrier()
    h2d, d2h, dup = C.c_double(), C.c_double(), (C.c_double * 2)()
    grok_b200._lib.check(grok_b200.lib.grkb_bench_link(256 << 20, 4, C.byref(h2d), C.byref(d2h), dup), "grkb_bench_link")
    barrier()
    clocks = sampler.stop()
    # ---- max over ranks
    keys = ["e2e_ms", "e2e16_ms"]
    vals = [main["tm"]["total_ms"]] + [main[k] for k in keys]
    if sub:
        vals += [sub["tm"]["total_ms"]] + [sub[k] for k in keys]
    link = [h2d.value, d2h.value, dup[0], dup[1]]
    all_clocks = [clocks]
    # (shapes that do not depend on what happened on a rank: grok_b200/shard.py, tested with gloo)
    leg_vals = {k: enc[k]["value"] for k in ENCODE_LEGS if enc and isinstance(enc.get(k), dict) and "value" in enc[k]}
    extra_ms = {"tiled_8k": tiled["ms"] if tiled else 0.0, "inflight4": main.get("inflight4_ms", 0.0)}
    vals, link, legs_red, extras_red = shard.reduce_bench_figures(world, COLLECTIVE_DEVICE if dist is not None else None, vals, link,
                                                                   leg_vals, ENCODE_LEGS, extra_ms, ("tiled_8k", "inflight4"))
    tiled_ms, tiled_ranks = extras_red["tiled_8k"]
    ms4_local, ms4_ranks = extras_red["inflight4"]
    if dist is not None:
        all_clocks = [None] * world
        dist.all_gather_object(all_clocks, clocks)
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return 0
    sm = [c["sm_mhz"] for c in all_clocks if c and c.get("sm_mhz")]
    clocks = {"sm_mhz": min(sm) if sm else None,
              "sm_max_mhz": max([c["sm_max_mhz"] for c in all_clocks if c and c.get("sm_max_mhz")] or [0]) or None,
              "reasons": sorted({r for c in all_clocks if c for r in c.get("reasons", [])}),
              "samples": sum(c.get("samples", 0) for c in all_clocks if c), "gpus_sampled": len(all_clocks),
              "note": "sm_mhz = lowest per-GPU median under load"}

    mp = W * H / 1e6
    peak, peak_kind = peaks()

    def figures(wl_, r, ms, e2e_ms, e2e16_ms):
        tm = r["tm"]
        d2h_b = int(NCOMP * W * H * 4 + 4)
        alg_last = 2 * 4 * W * H * NCOMP
        alg_all = sum(2 * 4 * (-(-W >> l)) * (-(-H >> l)) for l in range(NUMRES - 1)) * NCOMP
        ach = alg_last / (tm["dwt_last_ms"] * 1e-3) / 1e9 if tm["dwt_last_ms"] > 0 else 0.0
        ach_all = alg_all / (tm["dwt_ms"] * 1e-3) / 1e9 if tm["dwt_ms"] > 0 else 0.0
        traffic = None
        rp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
        if os.path.exists(rp) and not wl_["ht"]:
            traffic = json.load(open(rp)).get("idwt_last_level_dram_bytes")
        # what the link allows: the H2D and D2H of consecutive frames overlap (full duplex)
        t_link = max(r["h2d"] / (link[2] * 1e9), d2h_b / (link[3] * 1e9)) if link[2] > 0 and link[3] > 0 else None
        pcie_bound = world * mp / t_link if t_link else None
        e2e_v = world * mp / (e2e_ms * 1e-3)
        t1 = {"code_blocks_per_s": r["blocks"] / (tm["t1_ms"] * 1e-3) if tm["t1_ms"] > 0 else None,
              "bound": "warp-instruction issue: serial %s chain per code block (not HBM)"
                       % ("VLC/MEL/MagSgn" if wl_["ht"] else "MQ decision")}
        if not wl_["ht"] and os.path.exists(T1_PROFILE):
            prof = json.load(open(T1_PROFILE))
            t1.update({"decisions_per_frame": prof["decisions_per_frame"],
                       "decisions_per_s": prof["decisions_per_frame"] / (tm["t1_ms"] * 1e-3),
                       "warp_inst_per_decision": prof["warp_inst_per_decision"],
                       "issue_active_pct": prof["issue_active_pct"],
                       "from": prof["source"] + " (committed ncu capture, not measured in this run)"})
        return {
            "value": world * mp / (ms * 1e-3), "ms_per_step": ms,
            "e2e": {"value": e2e_v, "unit": "Mpixel/s", "frames_in_flight": r["K"], "h2d_bytes_per_step": r["h2d"],
                    "d2h_bytes_per_step": d2h_b, "ms_per_step": e2e_ms,
                    "boundary": "grkb_decompress_submit/_wait (C-ABI): codestream bytes in page-locked host memory in, "
                                "int32 planes out; own marker + packet parsing (T2) inside the timed region, like "
                                "the reference arm's",
                    "host_t2_ms_per_frame": r["t2_ms"]},
            "e2e_samples16": {"value": world * mp / (e2e16_ms * 1e-3), "unit": "Mpixel/s", "frames_in_flight": r["K"],
                              "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": int(NCOMP * W * H * 2 + 4),
                              "ms_per_step": e2e16_ms,
                              "note": "grkb_decompress_submit_packed: planes narrowed to uint16 on the device "
                                      "(SURVEY 8f N3, opt-in); `e2e` is the int32 grk_image contract"},
            "pcie_bound": {"value": pcie_bound, "unit": "Mpixel/s",
                           "what": "frames/s the link alone allows for this frame's H2D and D2H bytes, both "
                                   "directions busy at once, at the rates measured below"},
            "e2e_over_pcie_bound": (e2e_v / pcie_bound) if pcie_bound else None,
            "roofline": {"bound": "hbm", "kernel": "idwt_strip_kernel<%s,3 comps,fused %s> (last level)"
                                                    % (("K53", "RCT") if wl_["ht"] else ("K97", "ICT")),
                         "achieved": ach, "peak": peak, "peak_kind": peak_kind, "unit": "GB/s", "frac": ach / peak,
                         "traffic": traffic,
                         "traffic_note": "dram bytes of one ncu --set full capture under profiles/, not measured in "
                                         "this run" if traffic else None,
                         "algorithmic_bytes_per_launch": alg_last, "ms_per_launch": tm["dwt_last_ms"],
                         "whole_dwt": {"bytes": alg_all, "ms": tm["dwt_ms"], "achieved": ach_all, "frac": ach_all / peak,
                                       "what": "all %d levels of three components (SURVEY 8d: sum of 2*4*w_r*h_r), "
                                               "MCT fused" % (NUMRES - 1)}},
            "stages_ms": {("t1_ht_decode" if wl_["ht"] else "t1_mq_decode"): tm["t1_ms"],
                          "idwt_all_levels_plus_mct": tm["dwt_ms"], "idwt_last_level": tm["dwt_last_ms"]},
            "t1": t1,
            "parity": r.get("parity"),
            "detail": {"code_blocks": r["blocks"], "coded_bytes": r["coded_bytes"],
                       "max_passes_in_a_block": r["passes_max"], "kernel_launches_per_step": int(tm["launches"])},
        }

    f = figures(wl, main, vals[0], vals[1], vals[2])
    line = {"metric": "decode_mpixels_per_s", "value": f["value"], "unit": "Mpixel/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": f["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": wl["dtype"], "data": "synthetic",
            "config": config_dict(wl, world)}
    for k in ("e2e", "e2e_samples16", "pcie_bound", "e2e_over_pcie_bound", "roofline", "stages_ms", "t1", "parity",
              "detail"):
        line[k] = f[k]
    if ms4_local > 0 and ms4_ranks == world:
        ms4 = ms4_local
        line["value_4_frames_in_flight"] = {
            "value": world * W * H / 1e6 / (ms4 * 1e-3), "unit": "Mpixel/s", "ms_per_frame": ms4,
            "what": "device-resident like `value`, four decoders on four streams per GPU, wall clock over their "
                    "passes"}
    line["gpu_launches"] = int(main["tm"]["launches"]) * args.steps
    line["link"] = {"h2d_gbs": link[0], "d2h_gbs": link[1], "duplex_h2d_gbs": link[2], "duplex_d2h_gbs": link[3],
                    "what": "page-locked cudaMemcpyAsync, 256 MiB x 4, all %d ranks at once (lowest rank)" % world}
    line["detail"].update({"sharding": "one frame per GPU per step, no collective",
                           "host_threads_bound_to_gpu_numa_node": bool(numa_bound),
                           "l2": "working set per step (coded bytes + 2 x %d MB planes) exceeds the 126 MB L2"
                                 % (NCOMP * W * H * 4 // 1000000)})
    if sub:
        g = figures(second, sub, vals[3], vals[4], vals[5])
        line["ht"] = {"config": config_dict(second, world), "dtype": second["dtype"], "steps": max(5, args.steps // 2)}
        for k in ("value", "ms_per_step", "e2e", "e2e_samples16", "pcie_bound", "e2e_over_pcie_bound", "roofline",
                  "stages_ms", "t1", "parity", "detail"):
            line["ht"][k] = g[k]
    line["clocks"] = clocks
    # ---- N = 1 only: the reference's own plugin doors, and the CPU baseline beside it
    if world == 1 and not args.no_plugin:
        pl = plugin_legs(wl, args)
        if pl:
            line.update(pl)
        if second:
            pl = plugin_legs(second, args)
            if pl:
                line["ht"].update(pl)
    if enc is not None:
        for i_, k in enumerate(ENCODE_LEGS):
            if isinstance(enc.get(k), dict) and "value" in enc[k]:
                enc[k]["value_this_rank"] = enc[k]["value"]
                enc[k]["value"], enc[k]["ranks_counted"] = legs_red[k]
        enc["what"] = ("grkb_compress, frames sharded over the ranks (weak scaling): value = sum over ranks; host threads "
                       "per rank = cores / ranks")
        line["encode"] = enc
    if tiled and tiled_ranks == world:
        line["tiled_8k"] = {"value": T8K * T8K / 1e6 / (tiled_ms * 1e-3), "unit": "Mpixel/s", "ms_per_image": tiled_ms,
                            "scaling": "strong: the 64 tiles of one image over the ranks, tile t on rank t mod N",
                            "config": {"workload": "htj2k_decode_8192x8192_rgb16_53_rct_lossless_5lvl_cblk64_tiles1024",
                                       "tiles": 64, "coded_bytes": tiled["coded_bytes"]},
                            "boundary": "grkb_decompress_submit_tiles: codestream bytes in, int32 rectangles out, own T2 "
                                        "inside; max over ranks", "parity_ok": tiled.get("parity_ok"),
                            "d2h_bytes_per_image": T8K * T8K * NCOMP * 4}
    if world == 1 and not args.no_cpu_baseline and O.have_ref():
        threads = os.cpu_count() or 1
        v, dt = cpu_reference_decode(codestream(wl), 3, 1, threads)
        line["cpu_baseline"] = {"value": v, "unit": "Mpixel/s", "cores": threads, "kind": "reference",
                                "sample": "whole frame x3: stock grk decompress (public C API, memory stream, markers "
                                          "+ T2 + T1 + wavelet + MCT), %.0f ms/frame" % (dt * 1e3)}
        cli = cli_timings(wl)
        if cli:
            line["cpu_baseline"]["cli"] = cli
        if second:
            v, dt = cpu_reference_decode(codestream(second), 3, 1, threads)
            line["ht"]["cpu_baseline"] = {"value": v, "unit": "Mpixel/s", "cores": threads, "kind": "reference",
                                          "sample": "whole frame x3, %.0f ms/frame" % (dt * 1e3)}
            cli = cli_timings(second)
            if cli:
                line["ht"]["cpu_baseline"]["cli"] = cli
    emit(line)
    if dist is not None:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-plugin", action="store_true", help="skip the grk_plugin_* legs (N = 1 extras)")
    ap.add_argument("--no-ht", action="store_true", help="skip the configs[2] sub-line")
    ap.add_argument("--no-encode", action="store_true", help="skip the encode legs")
    ap.add_argument("--no-tiled", action="store_true", help="skip the configs[3] tile-sharded leg")
    ap.add_argument("--workload", default="j2k97", choices=["j2k97", "ht"])
    ap.add_argument("--inflight", type=int, default=None,
                    help="frames in flight on the e2e leg (default: 4 for the MQ workload, whose kernel tails "
                         "overlap; 2 for HT, which is bound by the bus)")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        return run_reference(args, wl)
    return run_ours(args, wl)


if __name__ == "__main__":
    sys.exit(main())
