// geometry.h -- internal: the code-block layout walk of geometry.cpp with the precinct structure
// the packet parser needs beside the flat block list.
#pragma once
#include "../../include/grk_b200.h"
#include <vector>

struct grkb_layout_res
{
	uint32_t gw, gh; // precinct grid of the resolution
	uint32_t pwe, phe; // precinct size exponents at this resolution
	uint32_t x0, y0, x1, y1; // resolution rectangle (canvas coordinates of the tile-component)
	uint32_t first_pb; // index of (band 0, precinct 0) in grkb_layout_index::pb
	uint32_t nbands;
};
struct grkb_layout_pb
{
	uint32_t first; // first block of this (band, precinct) in the flat list
	uint32_t cgw, cgh; // its code-block grid (the dimensions of its tag trees)
	uint32_t band_empty; // the whole band is empty: T2 skips it (t2/T2Decompress.cpp:286-288)
};
// pb is ordered resolution -> band -> precinct, like the flat list: the entry of (res, band, precinct)
// is pb[res.first_pb + band * gw * gh + precinct]
struct grkb_layout_index
{
	std::vector<grkb_layout_res> res;
	std::vector<grkb_layout_pb> pb;
};

int grkb_layout_walk(const grkb_layout_params* p, grkb_cblk_geom* out, uint32_t cap, uint32_t* count,
					 grkb_layout_index* idx);
