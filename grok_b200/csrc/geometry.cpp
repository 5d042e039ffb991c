// geometry.cpp -- code-block layout of a tile-component (host only, no CUDA): resolutions,
// bands, precincts and the code-block grid, in the order the reference walks them
// (resolution -> band -> precinct -> code block, row-major inside a precinct).
//
// Reproduces the integer geometry of
//   TileComponent::init             src/lib/jp2/tile/TileComponent.cpp:69-167
//   Resolution::init                src/lib/jp2/t1/Resolution.h:38-80
//   Subband::generatePrecinctBounds src/lib/jp2/t1/Subband.h:76-88
//   PrecinctImpl / getCodeBlockBounds  src/lib/jp2/t1/Precinct.h:44-69
//   band window placement           src/lib/jp2/tile/TileComponentWindowBuffer.h:188-221
// (ISO/IEC 15444-1 B.5-B.7) so that descriptor tables built here address the same samples as
// the reference's DecompressBlockExec / CompressBlockExec lists.
#include "geometry.h"
#include <algorithm>

namespace
{
struct R
{
	uint32_t x0, y0, x1, y1;
	uint32_t w() const { return x1 - x0; }
	uint32_t h() const { return y1 - y0; }
	bool empty() const { return x1 <= x0 || y1 <= y0; }
};
uint32_t cdp2(uint32_t a, uint32_t b)
{
	return (uint32_t)(((uint64_t)a + ((uint64_t)1 << b) - 1) >> b);
}
uint32_t fdp2(uint32_t a, uint32_t b)
{
	return a >> b;
}
R band_rect(R tc, uint32_t nd, uint32_t orient)
{
	if(nd == 0)
		return tc;
	const uint32_t bx = (orient & 1) ? (1u << (nd - 1)) : 0, by = (orient >> 1) ? (1u << (nd - 1)) : 0;
	R r;
	r.x0 = tc.x0 <= bx ? 0 : cdp2(tc.x0 - bx, nd);
	r.y0 = tc.y0 <= by ? 0 : cdp2(tc.y0 - by, nd);
	r.x1 = tc.x1 <= bx ? 0 : cdp2(tc.x1 - bx, nd);
	r.y1 = tc.y1 <= by ? 0 : cdp2(tc.y1 - by, nd);
	return r;
}
R isect(R a, R b)
{
	R r{std::max(a.x0, b.x0), std::max(a.y0, b.y0), std::min(a.x1, b.x1), std::min(a.y1, b.y1)};
	if(r.x1 < r.x0 || r.y1 < r.y0)
		r = R{0, 0, 0, 0};
	return r;
}
} // namespace

extern "C" int grkb_tile_layout(const grkb_layout_params* p, grkb_cblk_geom* out, uint32_t cap,
								 uint32_t* count)
{
	return grkb_layout_walk(p, out, cap, count, nullptr);
}

int grkb_layout_walk(const grkb_layout_params* p, grkb_cblk_geom* out, uint32_t cap, uint32_t* count,
					 grkb_layout_index* idx)
{
	if(!p || !count || p->numres == 0 || p->numres > 33 || p->x1 < p->x0 || p->y1 < p->y0)
		return -1;
	const R tc{p->x0, p->y0, p->x1, p->y1};
	uint32_t n = 0;
	for(uint32_t resno = 0; resno < p->numres; ++resno)
	{
		const R res = band_rect(tc, p->numres - 1 - resno, 0);
		const R lower = resno ? band_rect(tc, p->numres - resno, 0) : R{0, 0, 0, 0};
		uint32_t pwe = p->prec_w_exp[resno], phe = p->prec_h_exp[resno];
		// precinct partition of the resolution (canvas coordinates)
		const uint32_t px0 = fdp2(res.x0, pwe) << pwe, py0 = fdp2(res.y0, phe) << phe;
		const uint64_t px1 = (uint64_t)cdp2(res.x1, pwe) << pwe, py1 = (uint64_t)cdp2(res.y1, phe) << phe;
		const uint32_t gw = res.x0 == res.x1 ? 0 : (uint32_t)((px1 - px0) >> pwe);
		const uint32_t gh = res.y0 == res.y1 ? 0 : (uint32_t)((py1 - py0) >> phe);
		if(idx)
		{
			grkb_layout_res lr{};
			lr.gw = gw;
			lr.gh = gh;
			lr.pwe = pwe;
			lr.phe = phe;
			lr.x0 = res.x0;
			lr.y0 = res.y0;
			lr.x1 = res.x1;
			lr.y1 = res.y1;
			lr.first_pb = (uint32_t)idx->pb.size();
			lr.nbands = resno ? 3 : 1;
			idx->res.push_back(lr);
		}
		// in band coordinates the partition is halved for resno > 0
		uint32_t sx = px0, sy = py0;
		if(resno)
		{
			sx = cdp2(sx, 1);
			sy = cdp2(sy, 1);
			--pwe;
			--phe;
		}
		const uint32_t cwe = std::min<uint32_t>(p->cblkw_exp, pwe), che = std::min<uint32_t>(p->cblkh_exp, phe);
		const uint32_t nbands = resno ? 3 : 1;
		for(uint32_t bi = 0; bi < nbands; ++bi)
		{
			const uint32_t orient = resno ? bi + 1 : 0;
			const R band = band_rect(tc, resno ? p->numres - resno : p->numres - 1, orient);
			for(uint64_t pi = 0; pi < (uint64_t)gw * gh; ++pi)
			{
				const uint32_t ax = sx + (uint32_t)((pi % gw) << pwe), ay = sy + (uint32_t)((pi / gw) << phe);
				const R prec = isect(R{ax, ay, ax + (1u << pwe), ay + (1u << phe)}, band);
				const uint32_t gx0 = fdp2(prec.x0, cwe), gy0 = fdp2(prec.y0, che);
				const uint32_t gx1 = cdp2(prec.x1, cwe), gy1 = cdp2(prec.y1, che);
				const uint32_t cgw = gx1 - gx0, cgh = gy1 - gy0;
				if(idx)
					idx->pb.push_back(grkb_layout_pb{n, cgw, cgh, band.empty() ? 1u : 0u});
				for(uint64_t cb = 0; cb < (uint64_t)cgw * cgh; ++cb)
				{
					const uint32_t bx = (gx0 + (uint32_t)(cb % cgw)) << cwe, by = (gy0 + (uint32_t)(cb / cgw)) << che;
					const R blk = isect(R{bx, by, bx + (1u << cwe), by + (1u << che)}, prec);
					if(out && n < cap)
					{
						grkb_cblk_geom& g = out[n];
						g.cx0 = blk.x0;
						g.cy0 = blk.y0;
						g.w = (uint16_t)blk.w();
						g.h = (uint16_t)blk.h();
						g.x = blk.x0 - band.x0 + ((orient & 1) ? lower.w() : 0);
						g.y = blk.y0 - band.y0 + ((orient & 2) ? lower.h() : 0);
						g.res = (uint8_t)resno;
						g.band = (uint8_t)bi;
						g.orient = (uint8_t)orient;
						g.level = (uint8_t)(p->numres - 1 - resno);
						g.precinct = (uint32_t)pi;
						g.cblkno = (uint32_t)cb;
					}
					++n;
				}
			}
		}
	}
	*count = n;
	return (out && n > cap) ? -2 : 0;
}
