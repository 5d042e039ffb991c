// packet_order.h -- internal: the order in which the packets of a tile follow each other
// (ISO/IEC 15444-1 B.12; reference: t2/PacketIter.cpp:55-292 the five progressions, :306-355 which
// positions start a precinct, t2/PacketManager.cpp:150-176 POC bounds), shared by the packet parser
// (t2_parse.cpp) and the packet writer (t2_write.cpp).  Components are not sub-sampled here.
#pragma once
#include "codestream.h"
#include "geometry.h"
#include <algorithm>

namespace grkb_cs
{
// f(layer, resolution, component, precinct) is called once per packet position of every progression
// record, in order; it returns false to stop.  A packet visited by two records is the caller's to
// skip (the parser and the writer keep a "sent" map when there is more than one record).
template<class F>
static inline void for_each_packet(const std::vector<Poc>& progs, uint32_t numres, uint32_t nc, uint32_t numlayers,
								   const grkb_layout_index& idx, uint32_t tx0, uint32_t ty0, uint32_t tx1, uint32_t ty1,
								   F&& do_packet_fn)
{
	auto cdp2 = [](uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a + ((uint64_t)1 << b) - 1) >> b); };
	bool stop = false;
	auto do_packet = [&](uint32_t l, uint32_t r, uint32_t c, uint64_t p) {
		if(!stop && !do_packet_fn(l, r, c, p))
			stop = true;
	};
	// which (x, y) positions start precinct `(px, py)` of resolution r (B.12.1.3-5 with dx = dy = 1)
	auto precinct_at = [&](uint32_t resno, uint32_t x, uint32_t y, uint64_t& precinct) -> bool {
		const grkb_layout_res& R = idx.res[resno];
		const uint32_t lvl = numres - 1 - resno;
		if(R.gw == 0 || R.gh == 0 || R.x0 == R.x1 || R.y0 == R.y1)
			return false;
		const uint32_t rpx = R.pwe + lvl, rpy = R.phe + lvl;
		const bool okx = (rpx < 32 ? (x % (1u << rpx)) == 0 : x == 0) ||
						 (x == tx0 && (((uint64_t)R.x0 << lvl) % ((uint64_t)1 << rpx)) != 0);
		const bool oky = (rpy < 32 ? (y % (1u << rpy)) == 0 : y == 0) ||
						 (y == ty0 && (((uint64_t)R.y0 << lvl) % ((uint64_t)1 << rpy)) != 0);
		if(!okx || !oky)
			return false;
		const uint32_t gx = (cdp2(x, lvl) >> R.pwe) - (R.x0 >> R.pwe);
		const uint32_t gy = (cdp2(y, lvl) >> R.phe) - (R.y0 >> R.phe);
		if(gx >= R.gw || gy >= R.gh)
			return false;
		precinct = gx + (uint64_t)gy * R.gw;
		return true;
	};
	// smallest precinct pitch on the reference grid over the resolutions of a component range
	auto min_step = [&](uint32_t& dx, uint32_t& dy) {
		uint64_t mx = UINT32_MAX, my = UINT32_MAX;
		for(uint32_t r = 0; r < numres; ++r)
		{
			const uint32_t lvl = numres - 1 - r;
			const uint64_t sx = (uint64_t)1 << std::min<uint32_t>(idx.res[r].pwe + lvl, 40);
			const uint64_t sy = (uint64_t)1 << std::min<uint32_t>(idx.res[r].phe + lvl, 40);
			mx = std::min(mx, sx);
			my = std::min(my, sy);
		}
		dx = (uint32_t)std::min<uint64_t>(mx, 1u << 31);
		dy = (uint32_t)std::min<uint64_t>(my, 1u << 31);
	};
	uint32_t dxm = 1, dym = 1;
	min_step(dxm, dym);
	auto next_pos = [](uint32_t v, uint32_t d) -> uint64_t { return (uint64_t)v + d - (v % d); };

	for(const Poc& pr : progs)
	{
		if(stop)
			break;
		const uint32_t rS = pr.resS, rE = std::min<uint32_t>(pr.resE, numres);
		const uint32_t cS = pr.compS, cE = std::min<uint32_t>(pr.compE, nc);
		const uint32_t lE = std::min<uint32_t>(pr.layE, numlayers);
		switch(pr.prog)
		{
			case 0: // LRCP
				for(uint32_t l = 0; l < lE && !stop; ++l)
					for(uint32_t r = rS; r < rE && !stop; ++r)
						for(uint32_t c = cS; c < cE && !stop; ++c)
							for(uint64_t p = 0; p < (uint64_t)idx.res[r].gw * idx.res[r].gh && !stop; ++p)
								do_packet(l, r, c, p);
				break;
			case 1: // RLCP
				for(uint32_t r = rS; r < rE && !stop; ++r)
					for(uint32_t l = 0; l < lE && !stop; ++l)
						for(uint32_t c = cS; c < cE && !stop; ++c)
							for(uint64_t p = 0; p < (uint64_t)idx.res[r].gw * idx.res[r].gh && !stop; ++p)
								do_packet(l, r, c, p);
				break;
			case 2: // RPCL
				for(uint32_t r = rS; r < rE && !stop; ++r)
					for(uint64_t y = ty0; y < ty1 && !stop; y = next_pos((uint32_t)y, dym))
						for(uint64_t x = tx0; x < tx1 && !stop; x = next_pos((uint32_t)x, dxm))
							for(uint32_t c = cS; c < cE && !stop; ++c)
							{
								uint64_t p;
								if(!precinct_at(r, (uint32_t)x, (uint32_t)y, p))
									continue;
								for(uint32_t l = 0; l < lE && !stop; ++l)
									do_packet(l, r, c, p);
							}
				break;
			case 3: // PCRL
				for(uint64_t y = ty0; y < ty1 && !stop; y = next_pos((uint32_t)y, dym))
					for(uint64_t x = tx0; x < tx1 && !stop; x = next_pos((uint32_t)x, dxm))
						for(uint32_t c = cS; c < cE && !stop; ++c)
							for(uint32_t r = rS; r < rE && !stop; ++r)
							{
								uint64_t p;
								if(!precinct_at(r, (uint32_t)x, (uint32_t)y, p))
									continue;
								for(uint32_t l = 0; l < lE && !stop; ++l)
									do_packet(l, r, c, p);
							}
				break;
			default: // CPRL
				for(uint32_t c = cS; c < cE && !stop; ++c)
					for(uint64_t y = ty0; y < ty1 && !stop; y = next_pos((uint32_t)y, dym))
						for(uint64_t x = tx0; x < tx1 && !stop; x = next_pos((uint32_t)x, dxm))
							for(uint32_t r = rS; r < rE && !stop; ++r)
							{
								uint64_t p;
								if(!precinct_at(r, (uint32_t)x, (uint32_t)y, p))
									continue;
								for(uint32_t l = 0; l < lE && !stop; ++l)
									do_packet(l, r, c, p);
							}
				break;
		}
	}
}
} // namespace grkb_cs
