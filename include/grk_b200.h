/*
 * grk_b200.h -- C-ABI of the B200 (sm_100a) accelerator for Grok's JPEG 2000 tile-processing
 * hot path.  Plain C: pointers and sizes only, no C++ or torch types.
 *
 * Two layers are exported by libgrok_b200.so:
 *
 *   1. stage entry points (this header, section A): one call per reference stage, on HOST
 *      buffers laid out exactly like the reference's tile-component buffers (one int32 / float
 *      plane per component, Mallat layout in a single buffer, see below).  These are what a
 *      maintainer binds when replacing a single stage, and what the parity tests call.
 *   2. tile entry points (section B): whole-tile decode / encode with the block coder, the
 *      wavelet and the component transform chained on the device; used by the plugin
 *      (include/grk_b200_plugin.h) and by bench.py.
 *
 * Buffer layout ("tile buffer", reference: src/lib/jp2/tile/TileComponentWindowBuffer.h:188-221,
 * src/lib/jp2/cache/MemManager.cpp:35-43):  resolution r of a tile-component occupies the
 * top-left w_r x h_r samples of the plane; inside it LL is at (0,0), HL at (w_{r-1},0), LH at
 * (0,h_{r-1}) and HH at (w_{r-1},h_{r-1}).  9/7 data are IEEE single precision stored in the
 * same 32-bit cells.
 *
 * Every function returns 0 on success and a negative value on failure; nothing throws.
 * There is no CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef GRK_B200_H
#define GRK_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#pragma GCC visibility push(default)

#define GRKB_VERSION 1

/* library / device management -------------------------------------------------------------- */

/* returns GRKB_VERSION; never touches the device */
int grkb_version(void);
/* number of visible CUDA devices (0 if none), never fails */
int grkb_device_count(void);
/* selects the device used by the calling thread for all later calls */
int grkb_set_device(int device);
/* last error text of the calling thread (static storage) */
const char* grkb_last_error(void);

/* A. stage entry points, host buffers -------------------------------------------------------- */

/* Whole-tile wavelet transform of one tile-component, in place in `data`.
 * Replaces WaveletReverse::decompress (src/lib/jp2/transform/WaveletReverse.cpp:2232-2261,
 * whole-tile branch: decompress_tile_53 :802-877, decompress_tile_97 :1272-1347) and
 * WaveletFwdImpl::compress (src/lib/jp2/transform/WaveletFwd.cpp:618-629).
 *   data       int32 (reversible) or float (irreversible) plane, `stride` samples per row
 *   x0..y1     tile-component rectangle in canvas coordinates (origin decides lifting parity)
 *   numres     number of resolutions (decomposition levels + 1)
 *   reversible 1 = 5/3, 0 = 9/7
 *   forward    1 = analysis (encode), 0 = synthesis (decode) */
int grkb_dwt_host(void* data, uint32_t stride, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1,
				  uint32_t numres, int reversible, int forward);

/* Multi-component transform, in place on three planes of n samples.
 * mode 0: inverse RCT + DC shift + clamp  (mct::decompress_rev,   point_transform/mct.cpp:544-579)
 * mode 1: inverse ICT + DC shift + clamp  (mct::decompress_irrev, point_transform/mct.cpp:480-513)
 * mode 2: DC shift + forward RCT          (TileProcessor.cpp:506-535 + mct::compress_rev :583-587)
 * mode 3: DC shift + forward ICT          (TileProcessor.cpp:506-535 + mct::compress_irrev :591-596)
 * shift/lo/hi: per-component DC level shift and clamp range (lo/hi ignored for modes 2,3). */
int grkb_mct_host(void* c0, void* c1, void* c2, uint64_t n, int mode, const int32_t* shift,
				  const int32_t* lo, const int32_t* hi);

/* DC level shift of one plane, in place.
 * mode 0: x+shift, clamp            (mct::decompress_dc_shift_rev,   mct.cpp:515-539)
 * mode 1: rne(float x)+shift, clamp (mct::decompress_dc_shift_irrev, mct.cpp:450-475)
 * mode 2: x-shift                   (TileProcessor::dcLevelShiftCompress, TileProcessor.cpp:506-535) */
int grkb_dc_shift_host(void* c, uint64_t n, int mode, int32_t shift, int32_t lo, int32_t hi);

/* Decoder back end of one tile for `ncomp` components of identical geometry: inverse wavelet of
 * every component followed by the fused inverse MCT (mct != 0, ncomp == 3) or per-component DC
 * shift, exactly the tail of TileProcessor::decompressT1 + mctDecompress + dcLevelShiftDecompress
 * (src/lib/jp2/tile/TileProcessor.cpp:334-383, :457-504).
 *   planes[c]  in: dequantised coefficients (tile buffer layout); out: int32 samples */
int grkb_idwt_mct_host(void* const* planes, uint32_t ncomp, uint32_t stride, uint32_t x0,
					   uint32_t y0, uint32_t x1, uint32_t y1, uint32_t numres, int reversible,
					   int mct, const int32_t* shift, const int32_t* lo, const int32_t* hi);

/* Encoder front end, mirror of the above: DC shift, forward MCT (mct != 0, ncomp == 3) and
 * forward wavelet (TileProcessor::doCompress stages dcLevelShiftCompress, mct_encode, dwt_encode,
 * src/lib/jp2/tile/TileProcessor.cpp:202-232).
 *   planes[c]  in: int32 image samples; out: wavelet coefficients (tile buffer layout) */
int grkb_mct_fdwt_host(void* const* planes, uint32_t ncomp, uint32_t stride, uint32_t x0,
					   uint32_t y0, uint32_t x1, uint32_t y1, uint32_t numres, int reversible,
					   int mct, const int32_t* shift);

/* C. diagnostics ------------------------------------------------------------------------------ */

/* Times the device-resident wavelet back end (forward == 0: inverse DWT [+ fused inverse MCT])
 * or front end (forward == 1) on synthetic planes already in HBM, with CUDA events on the
 * launching stream: `warmup` untimed then `iters` timed passes; *ms_per_iter = mean pass time.
 * Used by bench.py for the DWT roofline figure. */
int grkb_bench_dwt(uint32_t w, uint32_t h, uint32_t numres, int reversible, int forward,
				   uint32_t ncomp, int mct, uint32_t iters, uint32_t warmup, float* ms_per_iter);

#pragma GCC visibility pop
#ifdef __cplusplus
}
#endif
#endif
