"""bench.py's assembly of the JSON line, with the GPU-facing pieces replaced by canned results (the control flow
of the extra legs -- one of them failed here -- and the keys of the contract); no GPU needed."""
import argparse
import contextlib
import io
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def test_line_assembly_with_a_failed_leg(monkeypatch):
    import bench
    import grok_b200

    class FakeLib:
        def grkb_device_count(self):
            return 1

        def grkb_set_device(self, d):
            return 0

        def grkb_bench_link(self, nbytes, iters, h2d, d2h, dup):
            h2d._obj.value, d2h._obj.value = 50.0, 51.0
            dup[0], dup[1] = 45.0, 46.0
            return 0

        def grkb_last_error(self):
            return b""

    fake = FakeLib()
    monkeypatch.setattr(grok_b200, "lib", fake)
    monkeypatch.setattr(grok_b200._lib, "lib", fake)

    def measure(wl, args, local, barrier, with_parity=True):
        r = {"blocks": 6708, "coded_bytes": 26000000, "t2_ms_single": 0.5, "passes_max": 58,
             "tm": dict(total_ms=15.3, t1_ms=15.2, dwt_ms=0.095, dwt_last_ms=0.045, launches=6),
             "e2e_ms": 12.7, "e2e16_ms": 12.5, "t2_ms": 0.48, "K": 4, "h2d": 26000000,
             "parity": {"checked": True, "max_abs_diff": 0}}
        if not wl["ht"]:
            r["inflight4_ms"] = 12.0
        return r

    def encode_legs(args, local=0, world=1):
        out = {k: {"value": 1700.0, "unit": "Mpixel/s"} for k in bench.ENCODE_LEGS[:2]}
        out[bench.ENCODE_LEGS[2]] = {"error": "this leg failed"}
        return out

    class Sampler:
        def __init__(self, local):
            pass

        def stop(self):
            return {"sm_mhz": 1965.0, "sm_max_mhz": 1965.0, "reasons": [], "samples": 10}

    monkeypatch.setattr(bench, "measure", measure)
    monkeypatch.setattr(bench, "codestream", lambda wl: None)
    monkeypatch.setattr(bench, "codestream8k", lambda: None)
    monkeypatch.setattr(bench, "tiled_leg", lambda *a: {"ms": 19.0, "coded_bytes": 277000000, "parity_ok": True})
    monkeypatch.setattr(bench, "encode_legs", encode_legs)
    monkeypatch.setattr(bench, "plugin_legs", lambda wl, args: {"e2e_plugin": {"value": 360.0}})
    monkeypatch.setattr(bench, "cpu_reference_decode", lambda *a: (31.0, 0.286))
    monkeypatch.setattr(bench, "cli_timings", lambda wl: None)
    monkeypatch.setattr(bench, "ClockSampler", Sampler)
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE"):
        monkeypatch.delenv(k, raising=False)
    args = argparse.Namespace(gpus=1, steps=20, warmup=3, no_plugin=False, no_ht=False, no_encode=False,
                              no_cpu_baseline=False, no_tiled=False, inflight=0, workload="j2k97", impl="ours")
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        assert bench.run_ours(args, bench.WORKLOADS["j2k97"]) == 0
    lines = [x for x in buf.getvalue().splitlines() if x.strip()]
    assert len(lines) == 1
    line = json.loads(lines[0])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches", "roofline", "clocks", "ht", "encode",
              "tiled_8k", "value_4_frames_in_flight", "cpu_baseline"):
        assert k in line, k
    assert line["n_gpus"] == 1 and abs(line["value"] - 4096 * 2160 / 1e6 / 15.3e-3) < 1e-6
    assert line["encode"][bench.ENCODE_LEGS[0]]["ranks_counted"] == 1
    assert "value" not in line["encode"][bench.ENCODE_LEGS[2]]


def _install_fakes(bench, grok_b200, rank):
    class FakeLib:
        def grkb_device_count(self):
            return 1

        def grkb_set_device(self, d):
            return 0

        def grkb_bind_host_to_device(self, d):
            return 0

        def grkb_bench_link(self, nbytes, iters, h2d, d2h, dup):
            h2d._obj.value, d2h._obj.value = 50.0 - rank, 51.0
            dup[0], dup[1] = 45.0, 46.0 - rank
            return 0

        def grkb_last_error(self):
            return b""

    fake = FakeLib()
    grok_b200.lib = fake
    grok_b200._lib.lib = fake

    def measure(wl, args, local, barrier, with_parity=True):
        barrier()
        return {"blocks": 6708, "coded_bytes": 26000000, "t2_ms_single": 0.5, "passes_max": 58,
                "tm": dict(total_ms=15.3 + rank, t1_ms=15.2, dwt_ms=0.095, dwt_last_ms=0.045, launches=6),
                "e2e_ms": 12.7 + rank, "e2e16_ms": 12.5, "t2_ms": 0.48, "K": 4, "h2d": 26000000,
                "parity": {"checked": True, "max_abs_diff": 0}}

    def encode_legs(args, local=0, world=1):
        if rank == 1:
            return {"error": "every leg failed on this rank"}
        return {k: {"value": 1700.0, "unit": "Mpixel/s"} for k in bench.ENCODE_LEGS}

    def tiled_leg(args, local, rank_, world, barrier):
        if rank_ == 1:
            raise RuntimeError("the tiled leg failed on this rank")
        return {"ms": 19.0, "coded_bytes": 277000000, "parity_ok": True}

    class Sampler:
        def __init__(self, local):
            pass

        def stop(self):
            return {"sm_mhz": 1965.0 - rank, "sm_max_mhz": 1965.0, "reasons": [], "samples": 10}

    bench.measure = measure
    bench.codestream = lambda wl: None
    bench.codestream8k = lambda: None
    bench.tiled_leg = tiled_leg
    bench.encode_legs = encode_legs
    bench.ClockSampler = Sampler
    bench.COLLECTIVE_BACKEND, bench.COLLECTIVE_DEVICE = "gloo", None


def _rank_main(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank), GRKB_NO_NUMA_BIND="1")
    # the line goes to file descriptor 1: give every rank a file there
    fd = os.open(os.path.join(out_dir, "stdout%d.txt" % rank), os.O_WRONLY | os.O_CREAT | os.O_TRUNC)
    os.dup2(fd, 1)
    import bench
    import grok_b200
    _install_fakes(bench, grok_b200, rank)
    args = argparse.Namespace(gpus=world, steps=20, warmup=3, no_plugin=False, no_ht=False, no_encode=False,
                              no_cpu_baseline=False, no_tiled=False, inflight=0, workload="j2k97", impl="ours")
    assert bench.run_ours(args, bench.WORKLOADS["j2k97"]) == 0
    sys.stdout.flush()


def test_two_ranks_with_legs_failing_on_one_of_them(tmp_path):
    """the N > 1 flow of bench.py over gloo: the ranks reduce, rank 0 alone prints the one line, and legs that
    failed on rank 1 only neither stall the run nor get into the line as if every rank had them"""
    import socket
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_rank_main, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    out0 = [x for x in open(tmp_path / "stdout0.txt").read().splitlines() if x.strip()]
    out1 = [x for x in open(tmp_path / "stdout1.txt").read().splitlines() if x.strip()]
    assert len(out0) == 1 and out1 == []
    line = json.loads(out0[0])
    assert line["n_gpus"] == 2
    assert abs(line["value"] - 2 * 4096 * 2160 / 1e6 / 16.3e-3) < 1e-6  # the slower rank's time
    assert abs(line["link"]["h2d_gbs"] - 49.0) < 1e-9 and line["clocks"]["sm_mhz"] == 1964.0
    leg = line["encode"]["dci4k_97_mq_cblk32_rate_cinema4k"]
    assert leg["value"] == 1700.0 and leg["ranks_counted"] == 1
    assert "tiled_8k" not in line and "value_4_frames_in_flight" not in line
