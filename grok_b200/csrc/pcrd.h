// pcrd.h -- internal (non-exported): rate allocation of one tile on the device (pcrd.cu), used by the
// codestream-level encoder (compress.cu) through the tile encoder (encode.cu).
#pragma once
#include "../../include/grk_b200.h"
#include <cuda_runtime.h>
#include <stdint.h>
#include <vector>

// one (band, precinct) of a packet that the packet header walks (non-empty bands only)
struct PcrdBand
{
	uint32_t leaf_first; // first entry of this band's code-block grid in leaf_block[], raster order
	uint32_t cgw, cgh; // the grid = the dimensions of its two tag trees
	uint32_t band_bps; // bit planes of the band (Mb): zero-bit-plane tree values are band_bps - numbps
	uint32_t node_first; // first node of this band's trees in the node scratch arrays
	uint32_t packet;
};
// one packet of the tile's single layer, in progression order
struct PcrdPacket
{
	uint32_t band_first, nbands;
	uint32_t leaf_first, nleaves; // all leaves of its bands, band after band
	uint32_t node_first, nnodes; // all tag-tree nodes of its bands
};

// host-side description of a tile geometry's packets, built once per geometry (compress.cu)
struct PcrdLayout
{
	std::vector<PcrdPacket> packets;
	std::vector<PcrdBand> bands;
	std::vector<int32_t> leaf_block; // index into the encoder's block table, -1 for a degenerate block
	std::vector<uint32_t> leaf_band; // band of every leaf
	uint32_t nnodes = 0;
};

struct PcrdResult
{
	uint32_t iterations; // bisection steps taken
	uint32_t error; // 1: a block has more bit planes than its band, 2: bit buffer too small
	double thresh; // the slope threshold of the layer
	uint64_t layer_bytes; // packets of the layer, headers and bodies (as the accounting counts them)
};

struct grkb_pcrd; // device tables + scratch of one geometry

int grkb_pcrd_create(const PcrdLayout& lay, uint32_t nblocks, grkb_pcrd** out);
void grkb_pcrd_destroy(grkb_pcrd* p);
// Enqueues the allocation of the tile whose block results are in coded / passes (device) on s and the
// copies of the chosen pass counts (npass_host, page-locked, nblocks entries) and of the result record
// (result_host, page-locked) back to the host.  max_bytes: the tile's byte budget; csty: Scod (SOP 2,
// EPH 4).
int grkb_pcrd_enqueue(grkb_pcrd* p, const grkb_coded* coded, const grkb_pass* passes, uint64_t max_bytes,
					  uint32_t csty, uint32_t* npass_host, PcrdResult* result_host, cudaStream_t s);

// The tile encoder runs the allocation right behind its block coders (same stream) for the next tile
// it is given: registered before grkb_encoder_submit, consumed by it (encode.cu).
struct grkb_encoder;
void grkb_encoder_set_pcrd(grkb_encoder* e, grkb_pcrd* p, uint64_t max_bytes, uint32_t csty, uint32_t* npass_host,
						   PcrdResult* result_host);
