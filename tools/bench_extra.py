#!/usr/bin/env python3
"""Measurements beside bench.py's headline line, for the other BASELINE configs (one JSON line each,
kept under profiles/):

  ht_encode   configs[2], encode half: 4096x2160 12-bit RGB -> HTJ2K lossless code blocks
              (DC shift + RCT + 5/3 + HT cleanup encode + packing), device-resident and end to end
              (H2D of the image, D2H of block table + coded bytes), stock CPU encoder beside it.
  tiled       configs[3]: 8192x8192 16-bit RGB, 1024x1024 tiles, HTJ2K lossless decode; the 64 tiles
              are independent units: round-robin over the ranks' GPUs (shard.shard_units) and,
              inside a GPU, over a few decoders kept in flight (grkb_decoder_submit / wait).

  mq_encode   configs[4]: a batch of DCI-4K frames (4096x2160 12-bit RGB, 9/7 + ICT, 32x32 code
              blocks, EBCOT/MQ, every pass coded with its rate and distortion for the host's PCRD),
              frames sharded over the ranks (frame f -> rank f mod N, no collective); stock CPU
              encoder at the cinema-4K byte budget beside it.

usage: python tools/bench_extra.py ht_encode|tiled|mq_encode [--steps K] [--inflight N] [--size S]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def synth(w, h, prec, seed):
    """low-pass 'natural-like' field (sum of random 2-D cosines, amplitude ~ 1/f) + noise"""
    rng = np.random.default_rng(seed)
    yy = np.arange(h, dtype=np.float32)[:, None]
    xx = np.arange(w, dtype=np.float32)[None, :]
    planes = []
    for c in range(3):
        f = np.zeros((h, w), np.float32)
        for k in range(6):
            fx, fy = rng.uniform(0.5, 24, size=2) / max(w, h)
            f += np.cos(2 * np.pi * (fx * xx + fy * yy) + rng.uniform(0, 6.28)) / (1 + k)
        f = (f - f.min()) / (f.max() - f.min())
        v = f * ((1 << prec) - 1) + rng.standard_normal((h, w)).astype(np.float32) * (1 << max(prec - 10, 0))
        planes.append(np.clip(np.rint(v), 0, (1 << prec) - 1).astype(np.int32))
    return planes


def ht_encode(args):
    import grok_b200
    from grok_b200 import encode as E, tile as T
    from oracle import oracle as O
    W, H, prec, nr = 4096, 2160, 12, 6
    grok_b200._lib.check(grok_b200.lib.grkb_set_device(0), "grkb_set_device")
    img = synth(W, H, prec, 4096)
    params = T.tile_params(0, 0, W, H, 3, nr, 1, 1, prec)
    blocks, geom, qcd = E.enc_block_table(0, 0, W, H, 3, nr, prec, 1, ht=True)
    enc = E.TileEncoder(params)
    pin = [T.PinnedArray((H, W), np.int32) for _ in range(3)]
    for p, a in zip(pin, img):
        p.array[:] = a
    src = [p.array for p in pin]
    n = len(blocks)
    out = (T.PinnedArray((n,), T.CODED_DTYPE).array, T.PinnedArray((n, T.ENC_MAX_PASSES), T.PASS_DTYPE).array,
           T.PinnedArray((W * H * 3 * 2,), np.uint8).array)
    coded, passes, data = enc.encode(src, blocks, out=out)
    assert not coded["failed"].any()
    # end to end: K encoders in flight (grkb_encoder_submit / wait), one frame each, so the H2D of
    # one frame overlaps the kernels and the D2H of another
    K = max(1, args.inflight)
    pblocks = T.PinnedArray((n,), T.ENC_CBLK_DTYPE)
    pblocks.array[:] = blocks
    encs = [enc] + [E.TileEncoder(params) for _ in range(K - 1)]
    outs = [out] + [(T.PinnedArray((n,), T.CODED_DTYPE).array,
                     T.PinnedArray((n, T.ENC_MAX_PASSES), T.PASS_DTYPE).array,
                     T.PinnedArray((W * H * 3 * 2,), np.uint8).array) for _ in range(K - 1)]

    def run(nframes):
        pend = [False] * K
        for f in range(nframes):
            k = f % K
            if pend[k]:
                encs[k].wait()
            if args.samples16:
                encs[k].submit_packed(src16.array, pblocks.array, outs[k])
            else:
                encs[k].submit(src, pblocks.array, outs[k])
            pend[k] = True
        for k in range(K):
            if pend[k]:
                encs[k].wait()

    if args.samples16:
        # row N3: the frame crosses the bus as uint16 planes, widened on the device
        src16 = T.PinnedArray((3, H, W), np.uint16)
        for c in range(3):
            src16.array[c] = img[c]
        encs[0].submit_packed(src16.array, pblocks.array, outs[0])
        c16, p16, d16 = encs[0].wait()
        assert np.array_equal(c16, coded) and bytes(d16) == bytes(data)
    run(2 * K)
    t0 = time.perf_counter()
    run(args.steps * K)
    e2e_ms = (time.perf_counter() - t0) * 1e3 / (args.steps * K)
    coded, passes, data = enc.encode(src, blocks, out=out)
    tm = enc.time_device(args.steps, 3)
    # round trip through the GPU decoder: lossless
    dblocks, dsegs = E.decode_table(blocks, coded, passes, geom, qcd, prec, 1, True)
    dec = T.TileDecoder(params)
    back = dec.decode(dblocks, dsegs, np.concatenate([data, np.zeros(16, np.uint8)]))
    lossless = dec.rejected == 0 and all(np.array_equal(a, b) for a, b in zip(back, img))
    cpu = None
    if O.have_ref() and not args.no_cpu:
        threads = os.cpu_count() or 1
        O.ref_codec_encode(img, prec, 0, ht=True, numres=nr, threads=threads)
        t0 = time.perf_counter()
        for _ in range(3):
            cs = O.ref_codec_encode(img, prec, 0, ht=True, numres=nr, threads=threads)
        dt = (time.perf_counter() - t0) / 3
        cpu = {"value": W * H / 1e6 / dt, "unit": "Mpixel/s", "cores": threads, "kind": "reference",
               "sample": "whole frame x3: stock grk compress (public C API, memory stream, T2 included), "
                         "%.0f ms/frame, %d bytes" % (dt * 1e3, len(cs))}
    mp = W * H / 1e6
    print(json.dumps({
        "metric": "encode_mpixels_per_s", "value": mp / (tm["total_ms"] * 1e-3), "unit": "Mpixel/s",
        "n_gpus": 1, "steps": args.steps, "ms_per_step": tm["total_ms"], "dtype": "i32", "data": "synthetic",
        "config": {"workload": "htj2k_encode_4096x2160_rgb12_53_rct_lossless_5lvl_cblk64_1tile",
                   "code_blocks": int(n), "coded_bytes": int(len(data))},
        "e2e": {"value": mp / (e2e_ms * 1e-3), "unit": "Mpixel/s", "ms_per_step": e2e_ms,
                "encoders_in_flight": K, "input_samples": "uint16 planes (N3)" if args.samples16 else "int32 planes",
                "h2d_bytes_per_step": int(W * H * 3 * (2 if args.samples16 else 4) + blocks.nbytes),
                "d2h_bytes_per_step": int(len(data) + coded.nbytes + n * int(coded["numpasses"].max()) * 16)},
        "stages_ms": {"dc_rct_fdwt": tm["dwt_ms"], "t1_ht_encode_pack": tm["t1_ms"]},
        "round_trip_lossless": bool(lossless), "cpu_baseline": cpu}), flush=True)


def tiled(args):
    import grok_b200
    from grok_b200 import encode as E, shard, tile as T
    from oracle import oracle as O
    rank, local, world = shard.dist_env()
    S, TS, prec, nr = args.size, 1024, 16, 6
    grok_b200._lib.check(grok_b200.lib.grkb_set_device(local), "grkb_set_device")
    img = synth(S, S, prec, 8192)
    tiles = shard.tile_grid(0, 0, S, S, 0, 0, TS, TS)
    mine = shard.shard_units(len(tiles), rank, world)
    params = T.tile_params(0, 0, TS, TS, 3, nr, 1, 1, prec)
    blocks, geom, qcd = E.enc_block_table(0, 0, TS, TS, 3, nr, prec, 1, ht=True)
    # set-up (untimed): this rank's tiles are coded with the GPU encoder
    enc = E.TileEncoder(params)
    coded_tiles = []
    for t in mine:
        x0, y0, x1, y1 = tiles[t]
        src = [np.ascontiguousarray(p[y0:y1, x0:x1]) for p in img]
        coded, passes, data = enc.encode(src, blocks)
        db, ds = E.decode_table(blocks, coded, passes, geom, qcd, prec, 1, True)
        pool = T.PinnedArray((len(data) + 16,), np.uint8)
        pool.array[:len(data)] = data
        pool.array[len(data):] = 0
        pb = T.PinnedArray((len(db),), T.CBLK_DTYPE)
        pb.array[:] = db
        ps = T.PinnedArray((len(ds),), T.SEG_DTYPE)
        ps.array[:] = ds
        coded_tiles.append((pb.array, ps.array, pool.array))
    enc.close()
    K = args.inflight
    decs = [T.TileDecoder(params, len(blocks), 1 << 22) for _ in range(K)]
    outs = [[T.PinnedArray((TS, TS), np.int32) for _ in range(3)] for _ in range(K)]
    outs16 = [T.PinnedArray((3, TS, TS), np.uint16) for _ in range(K)] if args.samples16 else None
    result = [np.zeros((S, S), np.int32) for _ in range(3)] if args.verify else None

    def planes_of(k):
        if outs16 is not None:  # row N3: the tile came back as uint16 planes
            return [outs16[k].array[c].astype(np.int32) for c in range(3)]
        return [p.array for p in outs[k]]

    def one_pass(verify):
        pending = [None] * K
        for i, t in enumerate(range(len(mine))):
            k = i % K
            if pending[k] is not None:
                decs[k].wait()
                if verify:
                    shard.composite(result, planes_of(k), tiles[mine[pending[k]]])
            b, s, d = coded_tiles[t]
            if outs16 is not None:
                decs[k].submit_packed(b, s, d, outs16[k].array)
            else:
                decs[k].submit(b, s, d, [p.array for p in outs[k]])
            pending[k] = t
        for k in range(K):
            if pending[k] is not None:
                decs[k].wait()
                if verify:
                    shard.composite(result, planes_of(k), tiles[mine[pending[k]]])

    one_pass(args.verify)
    ok = True
    if args.verify:
        for t in mine:
            x0, y0, x1, y1 = tiles[t]
            ok = ok and all(np.array_equal(result[c][y0:y1, x0:x1], img[c][y0:y1, x0:x1]) for c in range(3))
    one_pass(False)
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    shard.barrier(world)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        one_pass(False)
    ms = (time.perf_counter() - t0) * 1e3 / args.steps
    ms = shard.max_over_ranks(ms, world, device="cuda" if world > 1 else None)
    if rank != 0:
        return
    cpu = None
    if O.have_ref() and not args.no_cpu and world == 1:
        threads = os.cpu_count() or 1
        cs = O.ref_codec_encode(img, prec, 0, ht=True, numres=nr, tile=(TS, TS), threads=threads)
        t0 = time.perf_counter()
        O.ref_codec_decode(cs, 3, S, S, threads=threads, want_planes=False)
        dt = time.perf_counter() - t0
        cpu = {"value": S * S / 1e6 / dt, "unit": "Mpixel/s", "cores": threads, "kind": "reference",
               "sample": "whole image x1: stock grk decompress of the %d-tile codestream (T2 included), "
                         "%.0f ms" % (len(tiles), dt * 1e3)}
    nbytes = sum(int(d.size) for _, _, d in coded_tiles)
    print(json.dumps({
        "metric": "decode_mpixels_per_s", "value": S * S / 1e6 / (ms * 1e-3), "unit": "Mpixel/s",
        "n_gpus": world, "steps": args.steps, "ms_per_step": ms, "scaling": "strong", "dtype": "i32",
        "data": "synthetic", "timed": "end to end: H2D of descriptors + coded bytes, decode, D2H of %s planes"
                                     % ("uint16 (N3)" if args.samples16 else "int32"),
        "config": {"workload": "htj2k_decode_%dx%d_rgb16_53_rct_lossless_5lvl_cblk64_tiles1024" % (S, S),
                   "tiles": len(tiles), "tiles_per_rank": len(mine), "decoders_in_flight": K,
                   "coded_bytes_this_rank": nbytes, "sharding": "tile t -> rank t mod N, no collective"},
        "lossless": bool(ok) if args.verify else None, "cpu_baseline": cpu}), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def mq_encode(args):
    import grok_b200
    from grok_b200 import encode as E, shard, tile as T
    from oracle import oracle as O
    rank, local, world = shard.dist_env()
    W, H, prec, nr = 4096, 2160, 12, 6
    grok_b200._lib.check(grok_b200.lib.grkb_set_device(local), "grkb_set_device")
    nframes = args.frames
    mine = shard.shard_units(nframes, rank, world)
    base = synth(W + 8 * nframes, H, prec, 4096)  # slow pan of one field: frame f starts at column 8f
    params = T.tile_params(0, 0, W, H, 3, nr, 0, 1, prec)
    blocks, geom, qcd = E.enc_block_table(0, 0, W, H, 3, nr, prec, 0, ht=False, cblk_exp=(5, 5))
    enc = E.TileEncoder(params)
    pins = []
    for f in mine:
        if args.samples16:
            pin = T.PinnedArray((3, H, W), np.uint16)
            for c in range(3):
                pin.array[c] = base[c][:, 8 * f:8 * f + W]
            pins.append(pin.array)
            continue
        pin = [T.PinnedArray((H, W), np.int32) for _ in range(3)]
        for p, a in zip(pin, base):
            p.array[:] = a[:, 8 * f:8 * f + W]
        pins.append([p.array for p in pin])
    n = len(blocks)
    out = (T.PinnedArray((n,), T.CODED_DTYPE).array, T.PinnedArray((n, T.ENC_MAX_PASSES), T.PASS_DTYPE).array,
           T.PinnedArray((W * H * 3 * 2,), np.uint8).array)

    K = max(1, args.inflight)
    pblocks = T.PinnedArray((n,), T.ENC_CBLK_DTYPE)
    pblocks.array[:] = blocks
    encs = [enc] + [E.TileEncoder(params) for _ in range(K - 1)]
    outs = [out] + [(T.PinnedArray((n,), T.CODED_DTYPE).array,
                     T.PinnedArray((n, T.ENC_MAX_PASSES), T.PASS_DTYPE).array,
                     T.PinnedArray((W * H * 3 * 2,), np.uint8).array) for _ in range(K - 1)]

    host = {"submit": 0.0, "wait": 0.0}

    def one_pass_threads():
        # one host thread per encoder, frames k, k+K, ... each: an encoder takes its next frame
        # the moment its last one is back, whatever the others are doing (the calls release the GIL)
        import threading
        nbs = [0] * K

        def work(k):
            for src in pins[k::K]:
                encs[k].submit_packed(src, pblocks.array, outs[k], mct_norms=E.MCT_NORMS_IRREV, rate_control=True) \
                    if args.samples16 else \
                    encs[k].submit(src, pblocks.array, outs[k], mct_norms=E.MCT_NORMS_IRREV, rate_control=True)
                nbs[k] += len(encs[k].wait()[2])
        th = [threading.Thread(target=work, args=(k,)) for k in range(K)]
        for t in th:
            t.start()
        for t in th:
            t.join()
        return sum(nbs)

    def one_pass():
        if K > 1 and not args.single_thread:
            return one_pass_threads()
        # frames of this rank, K encoders in flight (grkb_encoder_submit / wait)
        nb, pend = 0, [False] * K
        for i, src in enumerate(pins):
            k = i % K
            ta = time.perf_counter()
            if pend[k]:
                nb += len(encs[k].wait()[2])
            tb = time.perf_counter()
            (encs[k].submit_packed if args.samples16 else encs[k].submit)(
                src, pblocks.array, outs[k], mct_norms=E.MCT_NORMS_IRREV, rate_control=True)
            pend[k] = True
            host["wait"] += tb - ta
            host["submit"] += time.perf_counter() - tb
        ta = time.perf_counter()
        for k in range(K):
            if pend[k]:
                nb += len(encs[k].wait()[2])
        host["wait"] += time.perf_counter() - ta
        return nb

    nb = one_pass()
    assert not out[0]["failed"].any()
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    shard.barrier(world)
    t0 = time.perf_counter()
    host["submit"] = host["wait"] = 0.0
    for _ in range(args.steps):
        one_pass()
    ms = (time.perf_counter() - t0) * 1e3 / args.steps
    host_ms = {k: v * 1e3 / args.steps / max(1, len(mine)) for k, v in host.items()}
    ms = shard.max_over_ranks(ms, world, device="cuda" if world > 1 else None)
    tm = enc.time_device(max(2, args.steps), 1)
    if rank != 0:
        return
    cpu = None
    if O.have_ref() and not args.no_cpu and world == 1:
        threads = os.cpu_count() or 1
        frame = [a[:, :W] for a in base]
        O.ref_set_rate(W * H * 3 * 1.5 / 1302083.0)  # cinema4K@24 frame budget, codestream/Profile.cpp:1006-1116
        try:
            O.ref_codec_encode(frame, prec, 1, ht=False, numres=nr, cblk=(32, 32), threads=threads)
            t0 = time.perf_counter()
            for _ in range(2):
                cs = O.ref_codec_encode(frame, prec, 1, ht=False, numres=nr, cblk=(32, 32), threads=threads)
            dt = (time.perf_counter() - t0) / 2
        finally:
            O.ref_set_rate(0)
        cpu = {"value": W * H / 1e6 / dt, "unit": "Mpixel/s", "cores": threads, "kind": "reference",
               "sample": "one frame x2: stock grk compress (public C API, memory stream, 9/7 + ICT, 32x32 blocks, "
                         "T1 + PCRD + T2 at the cinema-4K budget), %.0f ms/frame, %d bytes" % (dt * 1e3, len(cs))}
    mp = W * H / 1e6
    print(json.dumps({
        "metric": "encode_mpixels_per_s", "value": mp * world / (tm["total_ms"] * 1e-3), "unit": "Mpixel/s",
        "n_gpus": world, "steps": args.steps, "ms_per_frame_device": tm["total_ms"], "scaling": "strong",
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "j2k_batch_encode_dci4k_rgb12_97_ict_5lvl_cblk32_mq_all_passes",
                   "frames": nframes, "frames_per_rank": len(mine), "code_blocks": int(n),
                   "coded_bytes_per_frame": int(nb // max(1, len(mine))),
                   "sharding": "frame f -> rank f mod N, no collective",
                   "scope": "DC shift + ICT + 9/7 + quantise + EBCOT/MQ (all passes, rates, distortions) + packing; "
                            "PCRD and T2 stay on the host (SURVEY 8f N2)"},
        "e2e": {"value": mp * nframes / (ms * 1e-3), "unit": "Mpixel/s", "ms_per_batch": ms,
                "input_samples": "uint16 planes (N3)" if args.samples16 else "int32 planes",
                "h2d_bytes_per_frame": int(W * H * 3 * (2 if args.samples16 else 4) + blocks.nbytes),
                "encoders_in_flight": K,
                "host_threads": 1 if (K == 1 or args.single_thread) else K,
                "host_ms_per_frame_in_submit": host_ms["submit"], "host_ms_per_frame_in_wait": host_ms["wait"],
                "d2h_bytes_per_frame": int(nb // max(1, len(mine)) + out[0].nbytes +
                                           n * int(out[0]["numpasses"].max()) * 16)},
        "stages_ms": {"dc_ict_fdwt": tm["dwt_ms"], "t1_mq_encode_pack": tm["t1_ms"]},
        "cpu_baseline": cpu}), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def plugin_batch(args):
    """BASELINE configs[5] as the reference drives it: grk_plugin_batch_compress over a directory
    of DCI-4K PPM frames.  Our plugin reads the files and does everything before T2 on the GPU
    (worker pool, one device encoder per worker); the UNMODIFIED host callback does rate
    allocation + T2 + the file write on the host threads.  Timed wall clock, files included."""
    import shutil
    import tempfile
    from oracle import oracle as O
    W, H, prec, nr = 4096, 2160, 12, 6
    nframes = args.frames
    if not O.have_plugin():
        raise SystemExit("reference host / plugin not built")
    if args.inflight:
        os.environ["GRKB_BATCH_THREADS"] = str(args.inflight)
    base = synth(W + 8 * nframes, H, prec, 4096)
    root = tempfile.mkdtemp(prefix="grkb_batch_")
    try:
        indir, outdir = os.path.join(root, "in"), os.path.join(root, "out")
        os.makedirs(indir)
        os.makedirs(outdir)
        for f in range(nframes):
            O.write_pnm(os.path.join(indir, "f%04d.ppm" % f), [a[:, 8 * f:8 * f + W] for a in base], prec)
        O.ref_plugin_open(device=0, threads=args.host_threads)
        ratio = W * H * 3 * 1.5 / 1302083.0  # cinema4K@24 frame budget
        O.ref_set_rate(ratio)
        try:
            O.ref_plugin_batch_encode(indir, outdir, 1, numres=nr, cblk=(32, 32))  # warm-up
            t0 = time.perf_counter()
            for _ in range(args.steps):
                n = O.ref_plugin_batch_encode(indir, outdir, 1, numres=nr, cblk=(32, 32))
                assert n == nframes
            dt = (time.perf_counter() - t0) / args.steps
            sizes = [os.path.getsize(os.path.join(outdir, x)) for x in sorted(os.listdir(outdir))]
            cpu = None
            if not args.no_cpu:
                threads = os.cpu_count() or 1
                frame = [a[:, :W] for a in base]
                O.ref_codec_encode(frame, prec, 1, ht=False, numres=nr, cblk=(32, 32), threads=threads)
                t1 = time.perf_counter()
                cs = O.ref_codec_encode(frame, prec, 1, ht=False, numres=nr, cblk=(32, 32), threads=threads)
                ct = time.perf_counter() - t1
                same = np.array_equal(np.fromfile(os.path.join(outdir, "f0000.j2k"), dtype=np.uint8), cs)
                cpu = {"value": W * H / 1e6 / ct, "unit": "Mpixel/s", "cores": threads, "kind": "reference",
                       "sample": "one frame: stock grk compress, memory stream, %.0f ms; plugin output of that "
                                 "frame byte-identical: %s" % (ct * 1e3, same)}
        finally:
            O.ref_set_rate(0)
    finally:
        shutil.rmtree(root, ignore_errors=True)
    mp = W * H / 1e6
    print(json.dumps({
        "metric": "encode_mpixels_per_s", "value": mp * nframes / dt, "unit": "Mpixel/s", "n_gpus": 1,
        "steps": args.steps, "ms_per_frame": dt * 1e3 / nframes, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "grk_plugin_batch_compress_dci4k_rgb12_97_ict_5lvl_cblk32_cinema4k_rate",
                   "frames": nframes, "workers": os.environ.get("GRKB_BATCH_THREADS", "default"),
                   "host_library_threads": args.host_threads or "all",
                   "scope": "PPM read + GPU (DC shift, ICT, 9/7, quantise, EBCOT, packing) in the plugin; PCRD + T2 + "
                            "file write in the unmodified host callback; wall clock",
                   "bytes_per_frame": int(sum(sizes) // max(1, len(sizes)))},
        "cpu_baseline": cpu}), flush=True)


def plugin_batch_decode(args):
    """grk_plugin_init_batch_decompress + grk_plugin_batch_decompress over a directory of 4K HTJ2K
    (or, with --mq, 9/7 MQ) codestreams: host T2 + output callback of the UNMODIFIED library on the
    host threads, everything between in our plugin.  Wall clock, files included; the output
    callback dumps the int32 planes to files."""
    import shutil
    import tempfile
    from oracle import oracle as O
    W, H, prec, nr = 4096, 2160, 12, 6
    nframes = args.frames
    if not O.have_plugin():
        raise SystemExit("reference host / plugin not built")
    if args.inflight:
        os.environ["GRKB_BATCH_THREADS"] = str(args.inflight)
    base = synth(W + 8 * 4, H, prec, 4096)
    root = tempfile.mkdtemp(prefix="grkb_bdec_")
    try:
        indir, outdir = os.path.join(root, "in"), os.path.join(root, "out")
        os.makedirs(indir)
        os.makedirs(outdir)
        threads = os.cpu_count() or 1
        streams = [O.ref_codec_encode([a[:, 8 * f:8 * f + W] for a in base], prec, 1 if args.mq else 0,
                                      ht=not args.mq, numres=nr, threads=threads) for f in range(4)]
        for f in range(nframes):
            streams[f % 4].tofile(os.path.join(indir, "f%04d.j2k" % f))
        O.ref_plugin_open(device=0, threads=args.host_threads)
        O.ref_plugin_batch_decode(indir, outdir)  # warm-up
        t0 = time.perf_counter()
        for _ in range(args.steps):
            assert O.ref_plugin_batch_decode(indir, outdir) == nframes
        dt = (time.perf_counter() - t0) / args.steps
        want = O.ref_codec_decode(streams[1], 3, W, H, threads=threads)
        t1 = time.perf_counter()
        O.ref_codec_decode(streams[1], 3, W, H, threads=threads)
        ct = time.perf_counter() - t1
        got = O.read_plane_dump(os.path.join(outdir, "f0001.ppm"))
        same = all(np.array_equal(a, b) for a, b in zip(got, want))
    finally:
        shutil.rmtree(root, ignore_errors=True)
    mp = W * H / 1e6
    print(json.dumps({
        "metric": "decode_mpixels_per_s", "value": mp * nframes / dt, "unit": "Mpixel/s", "n_gpus": 1,
        "steps": args.steps, "ms_per_frame": dt * 1e3 / nframes, "dtype": "f32" if args.mq else "i32",
        "data": "synthetic",
        "config": {"workload": "grk_plugin_batch_decompress_4096x2160_rgb12_" +
                               ("97_ict_mq" if args.mq else "53_rct_ht_lossless") + "_5lvl_cblk64",
                   "frames": nframes, "workers": os.environ.get("GRKB_BATCH_THREADS", "default"),
                   "host_library_threads": args.host_threads or "all",
                   "coded_bytes_per_frame": int(len(streams[0])),
                   "scope": "file read + T2 in the unmodified host callback; T1, dequantisation, inverse wavelet, "
                            "inverse MCT in the plugin; int32 planes handed to the host's output callback, which "
                            "writes them to a file; wall clock"},
        "cpu_baseline": {"value": mp / ct, "unit": "Mpixel/s", "cores": threads, "kind": "reference",
                         "sample": "one frame: stock grk decompress, memory stream, %.0f ms; plugin samples of that "
                                   "frame identical: %s" % (ct * 1e3, same)}}), flush=True)


def compress(args):
    """The encode side of the codestream door (grkb_compress): frames in page-locked host memory in,
    complete codestreams out -- T2, markers and (for the rate-controlled DCI frames) the reference's
    PCRD bisection on our side.  One host thread and one compressor per frame in flight."""
    import threading
    from oracle import oracle as O
    from grok_b200 import codestream as CS
    from grok_b200 import tile as T
    W, H = 4096, 2160
    dci = not args.ht
    base = [p >> 4 for p in synth(W, H, 16, 4096)]
    pins = [T.PinnedArray((H, W), np.int32) for _ in range(3)]
    for p_, b in zip(pins, base):
        p_.array[:] = b
    img = [p_.array for p_ in pins]
    kw = dict(reversible=False, cblk=(32, 32), rate=W * H * 3 * 1.5 / 1302083.0) if dci else \
        dict(reversible=True, ht=True, cblk=(64, 64))
    nthreads = args.inflight or 8
    zs = [CS.Compressor(W, H, 3, 12, numres=6, **kw) for _ in range(nthreads)]
    outs = [z.encode(img) for z in zs]  # warm-up: encoders, pools
    per = max(1, args.frames // nthreads)

    def work(z):
        for _ in range(per):
            z.encode(img)

    t0 = time.perf_counter()
    th = [threading.Thread(target=work, args=(z,)) for z in zs]
    for t in th:
        t.start()
    for t in th:
        t.join()
    dt = time.perf_counter() - t0
    n = per * nthreads
    st = zs[0].stats()
    line = {"metric": "encode_mpixels_per_s", "value": W * H / 1e6 * n / dt, "unit": "Mpixel/s", "n_gpus": 1,
            "ms_per_frame": dt * 1e3 / n, "dtype": "f32" if dci else "i32", "data": "synthetic",
            "config": {"workload": "grkb_compress_4096x2160_rgb12_" + ("97_ict_mq_cblk32_rate_cinema4k_bytes" if dci
                                                                       else "53_rct_ht_lossless_cblk64"),
                       "frames": n, "host_threads": nthreads, "coded_bytes_per_frame": int(outs[0].size),
                       "scope": "int32 planes in page-locked host memory -> H2D -> DC shift, MCT, wavelet, quantiser, "
                                "block coders on the device -> D2H of streams + pass records -> " +
                                ("PCRD bisection + " if dci else "") + "packet headers, markers on the host; wall clock",
                       "one_frame_alone": {"device_ms": st["device_ms"], "host_t2_ms": st["t2_ms"]}}}
    if not args.no_cpu and O.have_ref():
        try:
            O.ref_set_rate(kw.get("rate", 0))
            t = time.perf_counter()
            ref = O.ref_codec_encode(base, 12, 0 if args.ht else 1, ht=bool(args.ht), numres=6, cblk=kw["cblk"],
                                     threads=os.cpu_count())
            ct = time.perf_counter() - t
        finally:
            O.ref_set_rate(0)
        line["cpu_baseline"] = {"value": W * H / 1e6 / ct, "unit": "Mpixel/s", "cores": os.cpu_count(), "kind": "reference",
                                "sample": "one frame, stock encoder on a memory stream, %.0f ms, %d bytes" % (ct * 1e3, ref.size)}
    print(json.dumps(line), flush=True)
    for z in zs:
        z.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["ht_encode", "tiled", "mq_encode", "plugin_batch", "plugin_batch_decode", "compress"])
    ap.add_argument("--ht", action="store_true", help="compress: HTJ2K lossless frames instead of rate-controlled DCI-4K")
    ap.add_argument("--frames", type=int, default=8)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--inflight", type=int, default=None)
    ap.add_argument("--size", type=int, default=8192)
    ap.add_argument("--verify", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--single-thread", action="store_true",
                    help="mq_encode: drive all encoders round-robin from one host thread instead of one thread each")
    ap.add_argument("--mq", action="store_true", help="plugin_batch_decode: 9/7 MQ frames instead of HT lossless")
    ap.add_argument("--host-threads", type=int, default=0,
                    help="plugin_batch: threads of the reference host's own pool (0 = all cores)")
    ap.add_argument("--samples16", action="store_true",
                    help="encode benches: send the frames as uint16 planes (grkb_encoder_submit_packed); tiled: receive "
                         "the tiles as uint16 planes (grkb_decoder_submit_packed)")
    args = ap.parse_args()
    if args.inflight is None and not args.what.startswith("plugin_batch"):
        args.inflight = 4 if args.what == "tiled" else 2
    {"ht_encode": ht_encode, "tiled": tiled, "mq_encode": mq_encode, "plugin_batch": plugin_batch,
     "plugin_batch_decode": plugin_batch_decode, "compress": compress}[args.what](args)


if __name__ == "__main__":
    main()
