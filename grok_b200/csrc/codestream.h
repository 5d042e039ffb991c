// codestream.h -- internal (non-exported) declarations of the codestream index and the packet
// parser (codestream.cpp, t2_parse.cpp) shared with the whole-image engine (decompress.cu).
#pragma once
#include "../../include/grk_b200.h"
#include <stddef.h>
#include <stdint.h>
#include <string>
#include <vector>

namespace grkb_cs
{
constexpr uint32_t MAX_RES = 33;
constexpr uint32_t MAX_BANDS = 3 * MAX_RES - 2;

struct Comp
{
	uint8_t prec = 0, sgnd = 0, dx = 1, dy = 1;
};
// COD / COC of one tile-component (SPcod / SPcoc)
struct Coding
{
	uint8_t numres = 0, cblkw_exp = 0, cblkh_exp = 0, cblk_sty = 0, reversible = 0;
	uint8_t pw[MAX_RES], ph[MAX_RES]; // precinct size exponents (15: none)
	uint8_t level = 0; // scope of the marker that set it: 1 main COD, 2 main COC, 3 tile COD, 4 tile COC
};
// QCD / QCC of one tile-component
struct Quant
{
	uint8_t qntsty = 0, numgbits = 0;
	uint8_t expn[MAX_BANDS];
	uint16_t mant[MAX_BANDS];
	uint8_t level = 0;
};
struct Poc
{
	uint8_t resS, resE, prog;
	uint16_t compS, compE, layE;
};
// everything a tile header can override
struct TileHeader
{
	uint8_t csty = 0; // Scod: bit 1 SOP, bit 2 EPH
	uint8_t prog = 0, mct = 0;
	uint16_t numlayers = 1;
	Coding cod[GRKB_MAX_COMPS];
	Quant q[GRKB_MAX_COMPS];
	uint8_t roishift[GRKB_MAX_COMPS] = {};
	std::vector<Poc> pocs;
};
struct Range
{
	size_t off, len;
};
struct Tile
{
	TileHeader hdr;
	bool seen = false;
	std::vector<Range> parts; // tile-part bodies (after SOD), in codestream order
};

struct Codestream
{
	const uint8_t* data = nullptr; // the caller's buffer: raw codestream or JP2 file
	size_t len = 0;
	size_t cs_off = 0; // SOC
	uint32_t x0 = 0, y0 = 0, x1 = 0, y1 = 0; // image area on the reference grid
	uint32_t tx0 = 0, ty0 = 0, tw = 0, th = 0, ntx = 0, nty = 0;
	uint32_t numcomps = 0;
	uint16_t rsiz = 0;
	bool cap = false; // CAP marker present (Part 15)
	Comp comps[GRKB_MAX_COMPS];
	TileHeader main;
	std::vector<Tile> tiles;
	bool truncated = false; // the last tile part runs past the end of the buffer
};

// Reads the main header and indexes every tile part.  err receives the reason on failure.
bool open(const uint8_t* data, size_t len, Codestream& cs, std::string& err);

struct TileTables
{
	grkb_tile_params params; // geometry of the (reduced) tile, shift / clamp per component
	std::vector<grkb_cblk> blocks;
	std::vector<grkb_seg> segs;
	// where the blocks' bytes are: either straight in the caller's buffer (every block's bytes are
	// one contiguous run of the codestream: single-layer streams) -- then data_off is relative to
	// src_off and src_len bytes go to the device as they are -- or gathered into `pool`
	bool zero_copy = true;
	size_t src_off = 0, src_len = 0;
	std::vector<uint8_t> pool;
	uint32_t packets = 0;
	bool truncated = false;
};

// T2 of one tile: walks its packets in progression order and decodes the packet headers into the
// decoder's descriptor tables.  reduce: highest resolutions to drop (cp_reduce); max_layers: 0 = all.
// window: x0, y0, x1, y1 in the coordinates of the REDUCED tile (canvas coordinates / 2^reduce), or
// null.  With a window only the code blocks its samples depend on are put in the tables.
bool parse_tile(const Codestream& cs, uint32_t tileno, uint32_t reduce, uint32_t max_layers, TileTables& out,
				std::string& err, const uint32_t* window = nullptr);
} // namespace grkb_cs
