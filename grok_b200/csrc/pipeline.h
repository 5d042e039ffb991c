// pipeline.h -- internal (non-exported) host-side declarations shared by the .cu files.
#pragma once
#include "dwt.cuh"

// kernels' host launchers
int grkb_launch_idwt_level(DwtLevelArgs a, int reversible, int ncomp, int fin, cudaStream_t s);
int grkb_launch_fdwt_level(DwtLevelArgs a, int reversible, int ncomp, int fin, cudaStream_t s);
int grkb_launch_mct(const MctArgs& p, int mode, cudaStream_t s);
int grkb_launch_dc(const MctArgs& p, int mode, cudaStream_t s);

// A group of 1 or 3 tile-components that share one geometry, resident on the device.
struct DwtGroup
{
	int ncomp;
	void* coeff[3]; // Mallat coefficient planes (T1 output / FDWT output)
	uint32_t coeff_stride;
	void* scratch[3][2]; // ping-pong planes for the intermediate resolutions
	uint32_t scratch_stride;
	void* image[3]; // full-resolution planes: IDWT destination / FDWT source
	uint32_t image_stride;
};

// bytes needed for one scratch plane of a tile-component with `numres` resolutions
size_t grkb_dwt_scratch_bytes(grkb_rect tc, uint32_t numres, uint32_t* stride_out);

// fin: 0 none, 1 DC shift per component, 2 MCT + DC shift (ncomp == 3)
int grkb_idwt_device(const DwtGroup& g, grkb_rect tc, uint32_t numres, int reversible, int fin,
					 const int32_t* shift, const int32_t* lo, const int32_t* hi, cudaStream_t s,
					 cudaEvent_t* ev_last = nullptr, uint32_t* launches = nullptr);
// ev_last: optional pair of events recorded around the last (largest) level's kernel
int grkb_fdwt_device(const DwtGroup& g, grkb_rect tc, uint32_t numres, int reversible, int fin,
					 const int32_t* shift, cudaStream_t s);

// narrow sample formats (samples.cu, row N3): int32 planes <-> 8/16-bit packed buffer, both on the device
int grkb_launch_pack_samples(void* const* planes, uint32_t numcomps, uint32_t stride, uint32_t W, uint32_t H,
							 uint32_t format, void* packed, size_t row_bytes, cudaStream_t s);
int grkb_launch_unpack_samples(const void* packed, size_t row_bytes, uint32_t format, int sgnd, void* const* planes,
							   uint32_t numcomps, uint32_t stride, uint32_t W, uint32_t H, cudaStream_t s);
size_t grkb_samples_min_row_bytes(uint32_t format, uint32_t W, uint32_t numcomps);
size_t grkb_samples_bytes(uint32_t format, uint32_t H, uint32_t numcomps, size_t row_bytes);

// grkb_decoder_submit_packed into a rectangle of a larger packed image (tile.cu)
struct grkb_decoder;
struct grkb_cblk;
struct grkb_seg;
int grkb_decoder_submit_packed_region(grkb_decoder* d, const grkb_cblk* blocks, uint32_t nblocks,
									  const grkb_seg* segs, uint32_t nsegs, const uint8_t* bytes, size_t nbytes,
									  void* out, uint32_t format, size_t row_bytes, size_t plane_bytes);

int grkb_decoder_submit_region(grkb_decoder* d, const grkb_cblk* blocks, uint32_t nblocks, const grkb_seg* segs,
							   uint32_t nsegs, const uint8_t* bytes, size_t nbytes, int32_t* const* planes,
							   uint32_t stride, uint32_t rx, uint32_t ry, uint32_t rw, uint32_t rh);

void grkb_set_error(const char* fmt, ...);
