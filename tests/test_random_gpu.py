"""Randomised parity sweep (fixed seeds): image sizes, component counts, precisions, tile and precinct
grids, code-block sizes and styles, progression orders, layers and rates drawn at random; the product
decodes the STOCK encoder's codestream to the STOCK decoder's samples, and encodes the image to the
STOCK encoder's bytes.  The hand-picked cases elsewhere cover what each feature needs; this covers the
combinations nobody thought of."""
import os

import numpy as np
import pytest

import codec_helpers as H
from grok_b200 import codestream as CS
from oracle import oracle as O

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not O.have_ref(), reason="compiled reference (oracle/_ref) not present")]

GROK_COMMENT = "Created by Grok     version 9.2.0"
PROGS = ["LRCP", "RLCP", "RPCL", "PCRL", "CPRL"]


def draw(seed):
    r = np.random.default_rng(seed)
    irr = int(r.integers(0, 2))
    ncomp = 3 if irr else int(r.choice([1, 3, 3, 4]))  # (the stock encoder writes nothing for 9/7 on one component)
    w, h = int(r.integers(33, 420)), int(r.integers(33, 330))
    prec = int(r.choice([8, 10, 12, 16]))
    ht = int(r.integers(0, 4) == 0) and not irr
    nr = int(r.integers(1, 7))
    cb = [(64, 64), (32, 32), (16, 16), (64, 16), (8, 32), (32, 64)][int(r.integers(0, 6))]
    tile = (0, 0) if r.integers(0, 2) else (int(r.choice([64, 96, 128, 200])), int(r.choice([64, 80, 128, 160])))
    prog = PROGS[int(r.integers(0, 5))]
    prc = None
    if r.integers(0, 2):
        p0 = int(r.choice([32, 64, 128, 256]))
        prc = [(p0, p0)] if r.integers(0, 2) else [(p0, p0), (max(p0 // 2, 16), max(p0 // 2, 16))]
    sop, eph = bool(r.integers(0, 3) == 0), bool(r.integers(0, 3) == 0)
    sty = 0 if ht else int(r.choice([0, 0, 0, 1, 2, 4, 8, 16, 32, 5, 63]))
    off = (0, 0)
    if tile == (0, 0) and r.integers(0, 3) == 0:
        off = (int(r.integers(1, 30)), int(r.integers(1, 30)))
    return dict(irr=irr, ncomp=ncomp, w=w, h=h, prec=prec, ht=ht, nr=nr, cb=cb, tile=tile, prog=prog, prc=prc, sop=sop,
                eph=eph, sty=sty, off=off, r=r)


def stock_opts(c, extra=""):
    o = ["prog=" + c["prog"]]
    if c["prc"]:
        o.append("prc=" + ":".join("%dx%d" % p for p in c["prc"]))
    if c["sop"]:
        o.append("sop")
    if c["eph"]:
        o.append("eph")
    if c["sty"]:
        o.append("sty=%d" % c["sty"])
    if c["off"] != (0, 0):
        o.append("off=%d:%d" % c["off"])
    if extra:
        o.append(extra)
    return ";".join(o)


def ours_kw(c):
    return dict(prog=c["prog"], precincts=c["prc"], sop=c["sop"], eph=c["eph"], cblk_sty=c["sty"], origin=c["off"])


def stock_encode(img, c, opts, rate=0.0):
    try:
        O.ref_set_rate(rate)
        O.ref_set_encode_options(opts)
        return O.ref_codec_encode(img, c["prec"], c["irr"], ht=bool(c["ht"]), numres=c["nr"], cblk=c["cb"], tile=c["tile"],
                                  threads=4)
    except RuntimeError:
        pytest.skip("the stock encoder refuses this combination")
    finally:
        O.ref_set_rate(0)
        O.ref_set_encode_options("")


def consistent(cs):
    """the tile parts chain up to EOC by their Psot (for a few combinations of all mode switches, layers and
    precincts the stock encoder writes a tile length that its own packets do not have)"""
    b = cs.tobytes()
    i = 2
    while i + 2 <= len(b):
        m = (b[i] << 8) | b[i + 1]
        if m == 0xFFD9:
            return True
        if i + 4 > len(b):
            return False
        if m == 0xFF90:
            psot = int.from_bytes(b[i + 6:i + 10], "big")
            if psot == 0:
                return True
            i += psot
        elif 0xFF30 <= m <= 0xFF3F:
            i += 2
        elif m >> 8 == 0xFF:
            i += 2 + ((b[i + 2] << 8) | b[i + 3])
        else:
            return False
    return False


# seeds whose parameters make the STOCK ENCODER abort (an assertion inside the reference library)
# (heap corruption -- free() / munmap_chunk(): invalid pointer, SIGSEGV -- found by running every seed's stock
# encode in a process of its own; all of them have TERMALL or LAZY blocks in precincts of 64x64 or less)
STOCK_ABORTS = {166, 215, 264, 1010, 1022, 1098, 2036, 2080, 2133}


@pytest.fixture(scope="module")
def dec(gpu):
    d = CS.Decompressor(0, slots=2)
    yield d
    d.close()


@pytest.mark.parametrize("seed", list(range(100, 200)) + list(range(1000, 1150)))
def test_random_decode(dec, seed):
    c = draw(seed)
    r = c["r"]
    img = H.synth_image(c["w"], c["h"], c["ncomp"], c["prec"], seed=seed)
    layers = ""
    if not c["ht"] and r.integers(0, 2):
        layers = "rates=" + ":".join(str(x) for x in sorted(r.choice([60, 40, 25, 12, 6, 3], size=int(r.integers(2, 4)),
                                                                     replace=False).tolist(), reverse=True))
    if seed in STOCK_ABORTS:
        pytest.skip("the stock encoder aborts on this combination")
    cs = stock_encode(img, c, stock_opts(c, layers))
    if cs.size < 200:
        pytest.skip("the stock encoder wrote no codestream for this combination")
    if not consistent(cs):
        pytest.skip("the stock encoder's tile lengths do not match its packets")
    reduce = int(r.integers(0, min(c["nr"], 3))) if r.integers(0, 3) == 0 else 0
    maxl = int(r.integers(1, 3)) if layers and r.integers(0, 2) else 0
    if layers and (c["sty"] & 5) and c["prog"] != "LRCP":
        # The stock decoder reads the headers of the packets it skips (layers beyond cp_layer, resolutions
        # beyond cp_reduce) but only readPacketData() advances a segment's pass count, so the next header of
        # a block with several codeword segments (LAZY, TERMALL) is split wrongly, the packet's length comes
        # out wrong and everything after it is garbage (t2/T2Decompress.cpp:63-110, :390-440).  Harmless
        # when the skipped packets come last (LRCP); otherwise there is nothing to compare with.
        reduce, maxl = 0, 0
    cso = CS.Codestream(cs)
    rw, rh = cso.image_size(reduce)
    cso.close()
    try:
        O.ref_set_decode_options(reduce, maxl)
        want = O.ref_codec_decode(cs, c["ncomp"], rw, rh, threads=4)
    finally:
        O.ref_set_decode_options(0, 0)
    got = dec.decode(cs, reduce=reduce, max_layers=maxl)
    assert dec.rejected == 0
    # the reference sizes a reduced image ceil(w / 2^r) x ceil(h / 2^r) (SIZMarker.cpp:54-57) although the
    # reduced canvas rectangle is ceil(x1 / 2^r) - ceil(x0 / 2^r) wide: with an odd origin there is one row or
    # column more than there are samples, and the stock decoder leaves it uninitialised
    x0, y0 = c["off"]
    up = lambda v: (v + (1 << reduce) - 1) >> reduce  # noqa: E731
    vw, vh = up(x0 + c["w"]) - up(x0), up(y0 + c["h"]) - up(y0)
    for k in range(c["ncomp"]):
        assert got[k].shape == want[k].shape
        assert np.array_equal(got[k][:vh, :vw], want[k][:vh, :vw]), \
            (seed, k, {x: c[x] for x in c if x != "r"}, layers, reduce, maxl)
    # a window of the image is that region of the full decode (only the tiles and code blocks it depends on
    # are decoded; compared with ourselves: the reference's windowed decode disagrees with its own full decode
    # when the image origin is not (0, 0))
    wx0, wy0 = int(r.integers(0, c["w"] - 8)), int(r.integers(0, c["h"] - 8))
    wx1, wy1 = int(r.integers(wx0 + 1, c["w"] + 1)), int(r.integers(wy0 + 1, c["h"] + 1))
    if up(x0 + wx1) == up(x0 + wx0) or up(y0 + wy1) == up(y0 + wy0):
        return  # (no sample of the window at this resolution: the door says so)
    win = dec.decode_window(cs, (wx0, wy0, wx1, wy1), reduce=reduce, max_layers=maxl)
    assert dec.rejected == 0
    ox, oy = up(x0), up(y0)
    for k in range(c["ncomp"]):
        crop = got[k][up(y0 + wy0) - oy:up(y0 + wy1) - oy, up(x0 + wx0) - ox:up(x0 + wx1) - ox]
        assert win[k].shape == crop.shape and np.array_equal(win[k], crop), (seed, k, (wx0, wy0, wx1, wy1), reduce, maxl)


@pytest.mark.parametrize("seed", list(range(200, 300)) + list(range(2000, 2150)))
def test_random_encode(gpu, seed):
    c = draw(seed)
    r = c["r"]
    img = H.synth_image(c["w"], c["h"], c["ncomp"], c["prec"], seed=seed)
    mode = int(r.integers(0, 3)) if not c["ht"] else 0  # 0 every pass, 1 one rate, 2 several layers
    kw = ours_kw(c)
    extra, rate = "", 0.0
    if mode == 1:
        rate = float(r.choice([4, 8, 15, 30]))
        kw["rate"] = rate
    elif mode == 2:
        rates = sorted(r.choice([60, 40, 25, 12, 6, 3], size=int(r.integers(2, 4)), replace=False).tolist(), reverse=True)
        extra = "rates=" + ":".join(str(x) for x in rates)
        kw["rates"] = rates
    if seed in STOCK_ABORTS:
        pytest.skip("the stock encoder aborts on this combination")
    want = stock_encode(img, c, stock_opts(c, extra), rate)
    if want.size < 200:
        pytest.skip("the stock encoder wrote no codestream for this combination")
    if not consistent(want):
        # (seed 210: the final packet simulation, whose total becomes Psot, also runs into the unchecked
        # header writes of DESIGN.md section 8 and counts one byte twice: Psot = tile length + 1, a stream the
        # stock decoder cannot read back.  Ours is the same stream with the right Psot.)
        pytest.skip("the stock encoder's tile lengths do not match its packets")
    z = CS.Compressor(c["w"], c["h"], c["ncomp"], c["prec"], reversible=not c["irr"], ht=bool(c["ht"]), numres=c["nr"],
                      cblk=c["cb"], tile=c["tile"], comment=GROK_COMMENT, **kw)
    got = z.encode(img)
    z.close()
    n = min(got.size, want.size)
    assert got.size == want.size and np.array_equal(got, want), \
        (seed, got.size, want.size, int(np.argmax(got[:n] != want[:n])) if n else -1, {x: c[x] for x in c if x != "r"}, mode)
