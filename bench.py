#!/usr/bin/env python
"""bench.py -- whole-frame JPEG 2000 decode throughput of the B200 tile decoder.

Workload (BASELINE.json configs[1]): 4096x2160 16-bit RGB, irreversible 9/7 + ICT, 5
decomposition levels, 64x64 code blocks, one tile, one layer, every coding pass kept.
A step = one decode of one frame: EBCOT/MQ block decoding with fused dequantisation, inverse
9/7 wavelet of the three components, inverse ICT + DC shift + clamp fused into the last level.
T2 (packet parsing) stays on the host in the reference's design and is outside the step.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

ours:       value = device-resident throughput (inputs already in HBM), e2e = through the
            C-ABI with pinned host buffers (H2D of descriptors + bytes, D2H of int32 planes).
reference:  the unmodified reference's own multithreaded CPU decoder (public C API, memory
            stream, oracle/_ref/libgrok_ref.so) on the same frame, all host threads.
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

W, H, NCOMP, PREC, NUMRES, CBLK = 4096, 2160, 3, 16, 6, 6
WORKLOAD = "j2k_decode_4096x2160_rgb16_97_ict_5lvl_cblk64_1tile"
SEED = 4096
HT = False  # --workload ht: BASELINE configs[2], 12-bit HTJ2K lossless (5/3 + RCT) decode


def select_workload(name):
    global PREC, WORKLOAD, HT
    if name == "ht":
        PREC, HT = 12, True
        WORKLOAD = "htj2k_decode_4096x2160_rgb12_53_rct_lossless_5lvl_cblk64_1tile"


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def cache_path(tag):
    key = hashlib.sha1(("%s-%d-%d-%d-%d-%d-%d-v4" % (tag, W, H, PREC, NUMRES, SEED, HT)).encode()).hexdigest()[:12]
    d = os.environ.get("GRKB_BENCH_CACHE", "/tmp")
    return os.path.join(d, "grkb_bench_%s_%s.npz" % (tag, key))


def synth_frame():
    import codec_helpers as Hh
    p = cache_path("img")
    if os.path.exists(p):
        return list(np.load(p)["img"])
    t = time.time()
    img = Hh.synth_image(W, H, NCOMP, PREC, SEED)
    np.savez(p + ".tmp.npz", img=np.stack(img))
    os.replace(p + ".tmp.npz", p)
    log("[bench] synthetic frame generated in %.1f s" % (time.time() - t))
    return img


def coded_tile(img):
    """descriptor table + byte pool of the frame (setup, untimed): forward path on the GPU
    stages (parity-tested), block coding on the host cores with the reference's T1 encoder
    when oracle/_ref is present, else the oracle port."""
    import codec_helpers as Hh
    import grok_b200
    from oracle import oracle as O
    p = cache_path("tile")
    if os.path.exists(p):
        z = np.load(p)
        return dict(blocks=z["blocks"], segs=z["segs"], data=z["data"])
    t = time.time()
    coefs = grok_b200.stages.mct_fdwt(img, 0, 0, NUMRES, 1 if HT else 0, 1, [1 << (PREC - 1)] * 3)
    tile = Hh.encode_tile(coefs, PREC, HT, NUMRES, (CBLK, CBLK), use_ref=O.have_ref(),
                          threads=os.cpu_count(), ht=HT)
    np.savez(p + ".tmp.npz", blocks=tile["blocks"], segs=tile["segs"], data=tile["data"])
    os.replace(p + ".tmp.npz", p)
    log("[bench] frame coded on the host in %.1f s: %d blocks, %.1f MB" %
        (time.time() - t, len(tile["blocks"]), tile["data"].size / 1e6))
    return tile


def codestream(img):
    from oracle import oracle as O
    p = cache_path("cs")
    if os.path.exists(p):
        return np.load(p)["cs"]
    t = time.time()
    cs = O.ref_codec_encode(img, PREC, 0 if HT else 1, ht=HT, numres=NUMRES,
                            cblk=(1 << CBLK, 1 << CBLK), threads=os.cpu_count())
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


def dist_env():
    from grok_b200 import shard
    return shard.dist_env()


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


def cpu_port_decode(tile):
    """no compiled reference on this box: the single-threaded oracle port on a bounded sample
    of the frame's code blocks (T1 + dequantisation only), scaled to the frame"""
    from oracle import oracle as O
    idx = np.arange(0, len(tile["blocks"]), max(1, len(tile["blocks"]) // 200))
    t = time.perf_counter()
    px = 0
    for i in idx:
        b = tile["blocks"][i]
        if b["numsegs"] == 0:
            continue
        sg = [(int(s["len"]), int(s["numpasses"])) for s in
              tile["segs"][b["seg_first"]:b["seg_first"] + b["numsegs"]]]
        raw = tile["data"][b["data_off"]:b["data_off"] + b["data_len"]].tobytes()
        O.orc_t1_decode(raw, sg, int(b["numbps"]), int(b["w"]), int(b["h"]), orient=int(b["orient"]))
        px += int(b["w"]) * int(b["h"])
    dt = time.perf_counter() - t
    return px / NCOMP / 1e6 / dt, len(idx)


def run_reference(args):
    rank, local, world = dist_env()
    if rank != 0:
        return 0
    from oracle import oracle as O
    threads = os.cpu_count() or 1
    base = {"impl": "reference", "metric": "decode_mpixels_per_s", "unit": "Mpixel/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "i32" if HT else "f32",
            "data": "synthetic", "gpu_launches": 0,
            "config": {"workload": WORKLOAD, "width": W, "height": H, "components": NCOMP,
                       "precision": PREC, "wavelet": "5/3" if HT else "9/7",
                       "mct": "RCT" if HT else "ICT", "levels": NUMRES - 1,
                       "block_coder": "HT" if HT else "MQ", "cblk": "64x64", "frames_per_step": 1}}
    img = synth_frame()
    if O.have_ref():
        cs = codestream(img)
        val, dt = cpu_reference_decode(cs, args.steps, args.warmup, threads)
        kind = "reference"
        sample = ("whole frame: stock grk decompress (public C API, memory stream, T2 included), "
                  "%d steps" % args.steps)
    else:
        tile = coded_tile(img)
        val, n = cpu_port_decode(tile)
        dt = W * H / 1e6 / val
        kind, threads = "port", 1
        sample = "oracle port, T1 decode of %d of %d code blocks, scaled" % (n, len(tile["blocks"]))
    base.update({"value": val, "ms_per_step": dt * 1e3,
                 "cpu_baseline": {"value": val, "unit": "Mpixel/s", "cores": threads, "kind": kind,
                                  "sample": sample},
                 "e2e": {"value": val, "unit": "Mpixel/s", "h2d_bytes_per_step": 0,
                         "d2h_bytes_per_step": 0}})
    print(json.dumps(base), flush=True)
    return 0


def run_ours(args):
    rank, local, world = dist_env()
    import grok_b200
    from grok_b200 import tile as T
    from oracle import oracle as O
    if grok_b200.lib.grkb_device_count() < 1:
        raise SystemExit("bench.py: no CUDA device visible; the product has no CPU fallback")
    grok_b200._lib.check(grok_b200.lib.grkb_set_device(local), "grkb_set_device")
    from grok_b200 import shard
    numa_bound = shard.bind_to_device(local) if world > 1 else False
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()

    # ---- setup (untimed): rank 0 builds the coded frame, the others read the cache
    if rank == 0:
        img = synth_frame()
        tile = coded_tile(img)
    barrier()
    if rank != 0:
        img = synth_frame()
        tile = coded_tile(img)
    blocks, segs, data = tile["blocks"], tile["segs"], tile["data"]
    params = T.tile_params(0, 0, W, H, NCOMP, NUMRES, 1 if HT else 0, 1, PREC)
    dec = T.TileDecoder(params, len(blocks), data.size)
    pin_data = T.PinnedArray((data.size,), np.uint8)
    pin_data.array[:] = data
    pin_blocks = T.PinnedArray((len(blocks),), T.CBLK_DTYPE)
    pin_blocks.array[:] = blocks
    pin_segs = T.PinnedArray((len(segs),), T.SEG_DTYPE)
    pin_segs.array[:] = segs
    blocks, segs = pin_blocks.array, pin_segs.array
    pin_out = [T.PinnedArray((H, W), np.int32) for _ in range(NCOMP)]
    out = [p.array for p in pin_out]

    # first pass also makes the frame resident for the device-timed loop
    dec.decode(blocks, segs, pin_data.array, out=out)
    if dec.rejected:
        raise SystemExit("bench.py: %d code blocks rejected" % dec.rejected)

    sampler = ClockSampler(local) if rank == 0 else None
    # ---- value: device-resident decode, CUDA events on the decoder's stream
    barrier()
    tm = dec.time_device(args.steps, args.warmup)
    barrier()
    ms = tm["total_ms"]
    # ---- e2e: through the C-ABI with pinned host buffers, every step's H2D (descriptors +
    # coded bytes) and D2H (three int32 planes) inside the timed region.  K decoders keep K
    # frames in flight (submit frame i, then wait for frame i-K+1), so one frame's D2H overlaps the
    # next frames' H2D + kernels; each step still moves and decodes a whole frame.
    K = max(1, args.inflight if args.inflight else (2 if HT else 4))
    decs = [(dec, out)]
    for _ in range(K - 1):
        pin_k = [T.PinnedArray((H, W), np.int32) for _ in range(NCOMP)]
        decs.append((T.TileDecoder(params, len(blocks), data.size), [p.array for p in pin_k]))

    def e2e_pass(n):
        for i in range(n):
            d, o = decs[i % K]
            if i >= K:
                d.wait()
            d.submit(blocks, segs, pin_data.array, o)
        for i in range(max(n - K, 0), n):
            decs[i % K][0].wait()

    e2e_pass(max(min(args.warmup, 3), K))
    barrier()
    t0 = time.perf_counter()
    e2e_pass(args.steps)
    e2e_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    barrier()
    if any(d.rejected for d, _ in decs) or not all(
            np.array_equal(a, b) for _, o in decs[1:] for a, b in zip(decs[0][1], o)):
        raise SystemExit("bench.py: pipelined decode disagrees between the decoders in flight")
    # ---- the same leg with the planes narrowed on the device (row N3, opt-in beyond the int32
    # grk_image contract): D2H of 16-bit samples instead of int32
    outs16 = [T.PinnedArray((NCOMP, H, W), np.uint16) for _ in range(K)]

    def e2e16_pass(n):
        for i in range(n):
            d = decs[i % K][0]
            if i >= K:
                d.wait()
            d.submit_packed(blocks, segs, pin_data.array, outs16[i % K].array)
        for i in range(max(n - K, 0), n):
            decs[i % K][0].wait()

    e2e16_pass(max(min(args.warmup, 3), K))
    barrier()
    t0 = time.perf_counter()
    e2e16_pass(args.steps)
    e2e16_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    barrier()
    if not all(np.array_equal(o.array[c], decs[0][1][c].astype(np.uint16)) for o in outs16 for c in range(NCOMP)):
        raise SystemExit("bench.py: 16-bit output differs from the int32 planes")
    clocks = sampler.stop() if sampler else None
    if dist is not None:
        import torch
        t = torch.tensor([ms, e2e_ms, e2e16_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_ms, e2e16_ms = float(t[0]), float(t[1]), float(t[2])
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return 0

    mp = W * H / 1e6
    peak, peak_kind = peaks()
    alg_last = 2 * 4 * W * H * NCOMP  # one read + one write of every sample of the last level
    traffic = None
    rp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(rp):
        traffic = json.load(open(rp)).get("idwt_last_level_dram_bytes")
    achieved = alg_last / (tm["dwt_last_ms"] * 1e-3) / 1e9 if tm["dwt_last_ms"] > 0 else 0.0
    h2d = int(blocks.nbytes + segs.nbytes + 4 * len(blocks) + data.size)
    d2h = int(NCOMP * W * H * 4 + 4)
    # ---- cpu baseline beside it (rank 0, N = 1 only): bounded, ~10-30 s of host work
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        if O.have_ref():
            cs = codestream(img)
            v, dt = cpu_reference_decode(cs, 3, 1, threads)
            cpu = {"value": v, "unit": "Mpixel/s", "cores": threads, "kind": "reference",
                   "sample": "whole frame x3: stock grk decompress (public C API, memory stream, "
                             "T2 included), %.0f ms/frame" % (dt * 1e3)}
        else:
            v, n = cpu_port_decode(tile)
            cpu = {"value": v, "unit": "Mpixel/s", "cores": 1, "kind": "port",
                   "sample": "oracle port, T1 decode of %d of %d code blocks, scaled" % (n, len(blocks))}
    line = {
        "metric": "decode_mpixels_per_s", "value": world * mp / (ms * 1e-3), "unit": "Mpixel/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "i32" if HT else "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "width": W, "height": H, "components": NCOMP,
                   "precision": PREC, "wavelet": "5/3" if HT else "9/7",
                   "mct": "RCT" if HT else "ICT", "levels": NUMRES - 1,
                   "block_coder": "HT" if HT else "MQ", "cblk": "64x64", "code_blocks": int(len(blocks)), "coded_bytes": int(data.size),
                   "frames_per_step": world, "sharding": "one frame per GPU per step, no collective",
                   "host_threads_bound_to_gpu_numa_node": bool(numa_bound),
                   "l2": "working set per step (coded bytes + 2 x %d MB planes) exceeds the 126 MB L2"
                         % (NCOMP * W * H * 4 // 1000000)},
        "e2e": {"value": world * mp / (e2e_ms * 1e-3), "unit": "Mpixel/s", "frames_in_flight": K,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms},
        "e2e_samples16": {"value": world * mp / (e2e16_ms * 1e-3), "unit": "Mpixel/s", "frames_in_flight": K,
                          "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": int(NCOMP * W * H * 2 + 4),
                          "ms_per_step": e2e16_ms,
                          "note": "grkb_decoder_submit_packed: planes narrowed to uint16 on the device "
                                  "(SURVEY 8f N3, opt-in); `e2e` above is the int32 grk_image contract"},
        "gpu_launches": int(tm["launches"]) * args.steps,
        "roofline": {"bound": "hbm", "kernel": "idwt_level_kernel<%s,3 comps,fused %s> (last level)"
                                                % (("K53", "RCT") if HT else ("K97", "ICT")),
                     "achieved": achieved, "peak": peak, "peak_kind": peak_kind, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic,
                     "algorithmic_bytes_per_launch": alg_last, "ms_per_launch": tm["dwt_last_ms"]},
        "stages_ms": {"t1_ht_decode" if HT else "t1_mq_decode": tm["t1_ms"], "idwt_all_levels_plus_mct": tm["dwt_ms"],
                      "idwt_last_level": tm["dwt_last_ms"]},
        "t1": {"code_blocks_per_s": len(blocks) / (tm["t1_ms"] * 1e-3) if tm["t1_ms"] > 0 else None,
               "bound": "serial %s rate per warp (not HBM)" % ("VLC/MEL/MagSgn" if HT else "MQ decision")},
        "clocks": clocks,
    }
    if cpu:
        line["cpu_baseline"] = cpu
    print(json.dumps(line), flush=True)
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
    ap.add_argument("--workload", default="j2k97", choices=["j2k97", "ht"])
    ap.add_argument("--inflight", type=int, default=None,
                    help="decoders (frames) in flight on the e2e leg (default: 4 for the MQ workload, whose kernel "
                         "tails overlap; 2 for HT, which is bound by the bus)")
    args = ap.parse_args()
    select_workload(args.workload)
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
