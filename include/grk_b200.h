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
/* Restricts the calling thread to the CPUs next to `device` (the local_cpulist of its PCI
 * function), so that page-locked buffers allocated afterwards live on the GPU's own NUMA node.
 * For hosts that run one thread or process per GPU (plugin deviceId = -1, torchrun ranks).
 * Returns 0 when the affinity was narrowed, 1 when the box exposes no topology (nothing changed),
 * negative on a CUDA error. */
int grkb_bind_host_to_device(int device);
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

/* Array-based ("custom matrix") multi-component transform of Part 2, in place on `numcomps` planes
 * of n samples (row M7).  matrix: numcomps x numcomps, row-major (output component j = row j).
 * forward 0: mct::decompress_custom (point_transform/mct.cpp:660-686): planes hold IEEE single
 *            precision samples, out[j] = sum_k M[j][k] * in[k] with separately rounded products and
 *            sums in k order; the per-component DC shift follows as a separate stage
 *            (grkb_dc_shift_host), as in TileProcessor::dcLevelShiftDecompress (:490-504)
 * forward 1: mct::compress_custom (:622-658): int32 planes, the matrix truncated to 13-bit fixed
 *            point, products rounded by fix_mul (util/grk_intmath.h:66-79); like the reference,
 *            every output component uses the matrix's FIRST row (its pointer is reset per
 *            component, :642 -- a reference defect reproduced for parity) */
int grkb_mct_custom_host(void* const* planes, uint32_t numcomps, uint64_t n, const float* matrix, int forward);

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

/* B. tile entry points --------------------------------------------------------------------- */

/* One code block of a tile, as the reference's schedulers describe it
 * (DecompressBlockExec / CompressBlockExec, src/lib/jp2/t1/BlockExec.h:22-84, filled by
 * T1DecompressScheduler::prepareScheduleDecompress, t1/T1DecompressScheduler.cpp:27-75).
 * Coordinates are relative to the tile-component buffer (Mallat layout above), i.e. the band
 * offset is already included: x = cblk.x0 - band.x0 (+ w_{r-1} for HL/HH), same for y. */
typedef struct grkb_cblk
{
	uint32_t x, y; /* top-left sample in the tile-component buffer */
	uint16_t w, h; /* block size (w*h <= 4096, both <= 1024) */
	uint8_t comp; /* component index */
	uint8_t orient; /* 0 LL, 1 HL, 2 LH, 3 HH */
	uint8_t numbps; /* MQ: cblk->numbps (magnitude bit-planes to decode / produced by encode) */
	uint8_t cblk_sty; /* GRKB_CBLKSTY_* (mode switches; HT blocks: GRKB_CBLKSTY_HT) */
	uint8_t roishift; /* region-of-interest up-shift (decode) */
	uint8_t k_msbs; /* HT: band->numbps - cblk->numbps (missing MSBs), BlockExec.h:40 */
	uint16_t numsegs; /* decode: codeword segments of this block (0: block not coded) */
	uint32_t seg_first; /* decode: index of the first segment in the segment table */
	uint32_t data_off; /* byte offset of the block's first byte in the byte pool */
	uint32_t data_len; /* decode: total bytes; encode: capacity on input, bytes on output */
	float scale; /* decode 9/7: MQ stepsize/2, HT stepsize (PostDecompressFilters.h:66-89,161-177);
				  * encode 9/7: band stepsize (T1Part1.cpp:75-76) */
} grkb_cblk;

/* HT blocks (GRKB_CBLKSTY_HT): one segment with one pass is a cleanup-only block (all that
 * Grok's own host produces, T1HT.cpp:171-173).  Two segments are cleanup | SigProp [+ MagRef]:
 * segment 0 holds the cleanup pass (numpasses 1, ojph_decode_codeblock's lengths1), segment 1 the
 * refinement passes (numpasses 1 or 2, lengths2), ojph_block_decoder.cpp:989-1057.  More than one
 * pass without a second segment is rejected like there (:1014-1020). */
typedef struct grkb_seg
{
	uint32_t len; /* bytes of this codeword segment */
	uint32_t numpasses; /* coding passes in it */
} grkb_seg;

#define GRKB_CBLKSTY_LAZY 0x01
#define GRKB_CBLKSTY_RESET 0x02
#define GRKB_CBLKSTY_TERMALL 0x04
#define GRKB_CBLKSTY_VSC 0x08
#define GRKB_CBLKSTY_PTERM 0x10
#define GRKB_CBLKSTY_SEGSYM 0x20
#define GRKB_CBLKSTY_HT 0x40

#define GRKB_MAX_COMPS 4

/* Geometry and sample format of one tile (all components share the tile-component rectangle:
 * no sub-sampling in this round). */
typedef struct grkb_tile_params
{
	uint32_t x0, y0, x1, y1; /* tile-component rectangle, canvas coordinates */
	uint32_t numcomps; /* 1..GRKB_MAX_COMPS */
	uint32_t numres; /* resolutions = decomposition levels + 1 */
	int32_t reversible; /* 1: 5/3 + RCT, 0: 9/7 + ICT */
	int32_t mct; /* 1: components 0..2 go through the RCT/ICT */
	int32_t shift[GRKB_MAX_COMPS]; /* DC level shift (tccp->m_dc_level_shift) */
	int32_t lo[GRKB_MAX_COMPS], hi[GRKB_MAX_COMPS]; /* clamp range (mct.cpp:556-571) */
} grkb_tile_params;

/* Code-block layout of one tile-component, in the reference's walking order (resolution ->
 * band -> precinct -> code block; TileComponent::init, tile/TileComponent.cpp:69-167,
 * t1/Resolution.h:38-80, t1/Precinct.h:44-69).  Host-only helper: no device involved. */
typedef struct grkb_layout_params
{
	uint32_t x0, y0, x1, y1; /* tile-component rectangle, canvas coordinates */
	uint32_t numres;
	uint32_t cblkw_exp, cblkh_exp; /* log2 nominal code-block size */
	uint32_t prec_w_exp[33], prec_h_exp[33]; /* log2 precinct size per resolution (15 = none) */
} grkb_layout_params;

typedef struct grkb_cblk_geom
{
	uint32_t x, y; /* position in the tile-component buffer (Mallat layout) */
	uint32_t cx0, cy0; /* position in band coordinates (cblk->x0, cblk->y0) */
	uint16_t w, h; /* may be 0 for degenerate blocks the reference also enumerates */
	uint8_t res, band, orient, level; /* resolution, band index in it, orientation, decomposition level */
	uint32_t precinct, cblkno;
} grkb_cblk_geom;

/* Fills out[0..cap) and *count (total blocks; call with out == NULL to size).  Returns 0, or -2
 * if cap was too small. */
int grkb_tile_layout(const grkb_layout_params* p, grkb_cblk_geom* out, uint32_t cap,
					 uint32_t* count);

typedef struct grkb_decoder grkb_decoder;

/* Creates a tile decoder for one geometry on the calling thread's device: device planes, pinned
 * staging and a stream.  max_blocks / max_bytes bound the per-tile descriptor table and byte
 * pool (they grow on demand if exceeded). */
int grkb_decoder_create(const grkb_tile_params* p, uint32_t max_blocks, size_t max_bytes,
						grkb_decoder** out);
void grkb_decoder_destroy(grkb_decoder* d);

/* Whole-tile decode, the device side of TileProcessor::decompressT1 + mctDecompress
 * (src/lib/jp2/tile/TileProcessor.cpp:334-383, :457-504) for all components at once:
 * H2D of descriptors and code-block bytes, T1 (EBCOT/MQ or HT per block) with dequantisation
 * fused into the store, inverse wavelet of every component, inverse MCT + DC shift + clamp fused
 * into the last level, D2H of the int32 planes.
 *   blocks/segs/bytes  host memory (bytes may be pinned; it is staged through pinned memory
 *                      otherwise)
 *   planes[c]          host destination, `stride` samples per row, (x1-x0) x (y1-y0) samples;
 *                      NULL planes: results stay on the device (see grkb_decoder_download)
 * status_out (may be NULL): number of blocks the coder rejected (they decode as zeros, the
 * reference fails the tile: T1.cpp:1373-1377, ojph_block_decoder.cpp:1029-1043). */
int grkb_decoder_decode(grkb_decoder* d, const grkb_cblk* blocks, uint32_t nblocks,
						const grkb_seg* segs, uint32_t nsegs, const uint8_t* bytes, size_t nbytes,
						int32_t* const* planes, uint32_t stride, uint32_t* status_out);

/* The same in two halves, for keeping several tiles/frames in flight (one decoder each):
 * grkb_decoder_submit enqueues the copies and kernels on the decoder's stream and returns;
 * grkb_decoder_wait blocks until they are done.  blocks/segs/bytes/planes must stay valid and
 * untouched until the wait returns (pinned memory makes the copies truly asynchronous), and a
 * decoder takes one submit at a time. */
int grkb_decoder_submit(grkb_decoder* d, const grkb_cblk* blocks, uint32_t nblocks,
						const grkb_seg* segs, uint32_t nsegs, const uint8_t* bytes, size_t nbytes,
						int32_t* const* planes, uint32_t stride);
int grkb_decoder_wait(grkb_decoder* d, uint32_t* status_out);

/* Narrow sample formats at the bus (opt-in, beyond the int32 grk_image
 * contract).  The reference narrows int32 planes on the host whenever samples leave the library
 * (planarToInterleaved, src/bin/common/convert.cpp:397-405; the 8/16-bit file writers) and widens
 * file samples on the way in (convert_16u32s_C1R, convert.cpp:547-556); here that runs on the
 * device and the bus carries 1 or 2 bytes per sample.  Samples are in host byte order: the low
 * 16 / 8 bits of the int32 (two's complement for signed components).  Planar buffers hold
 * numcomps planes of H rows back to back; interleaved ones H rows of W pixels of numcomps
 * samples; rows are row_bytes apart (>= the samples of a row; a multiple of 8 with an 8-byte
 * aligned base keeps the device side on vector accesses). */
enum
{
	GRKB_SAMPLES_32 = 0, /* int32 planes: the grk_image form */
	GRKB_SAMPLES_16 = 1,
	GRKB_SAMPLES_16_INTERLEAVED = 2,
	GRKB_SAMPLES_8 = 3,
	GRKB_SAMPLES_8_INTERLEAVED = 4,
	GRKB_SAMPLES_BIG_ENDIAN = 0x10 /* OR-ed to a 16-bit format: samples as PNM files store them */
};
/* grkb_decoder_submit with the result narrowed on the device: `out` receives
 * grkb_samples_size(format, ...) bytes.  Precision above the format's width is refused. */
int grkb_decoder_submit_packed(grkb_decoder* d, const grkb_cblk* blocks, uint32_t nblocks,
							   const grkb_seg* segs, uint32_t nsegs, const uint8_t* bytes, size_t nbytes,
							   void* out, uint32_t format, size_t row_bytes);
/* bytes of a packed buffer: H rows (x numcomps planes when planar) of row_bytes */
size_t grkb_samples_size(uint32_t format, uint32_t H, uint32_t numcomps, size_t row_bytes);

/* Re-runs the device stages (T1, wavelet, MCT) of the last tile given to grkb_decoder_decode,
 * inputs already resident in HBM, no host copies: `iters` passes after `warmup`, timed with CUDA
 * events on the decoder's stream.  ms[0] = mean whole pass, ms[1] = mean T1 kernels,
 * ms[2] = mean of all wavelet levels, ms[3] = mean of the last (largest) wavelet level kernel.
 * launches_out (may be NULL): kernels launched per pass. */
int grkb_decoder_time_device(grkb_decoder* d, uint32_t iters, uint32_t warmup, float ms[4],
							 uint32_t* launches_out);

/* Copies the decoded planes of the last pass to host memory. */
int grkb_decoder_download(grkb_decoder* d, int32_t* const* planes, uint32_t stride);

/* Stage-level door for the block decoder alone (parity tests): decodes `nblocks` blocks into a
 * dense plane set exactly as T1Part1::decompress / T1HT::decompress + postProcess would
 * (t1/t1_part1/T1Part1.cpp:133-161, t1/t1_ht/T1HT.cpp:134-187, tile/TileComponent.cpp:333-430).
 * planes[c]: host int32/float planes of plane_h rows x stride samples, written only inside the
 * block rectangles.  reversible selects the /2 (1) or x scale (0) dequantiser. */
int grkb_t1_decode_host(const grkb_cblk* blocks, uint32_t nblocks, const grkb_seg* segs,
						uint32_t nsegs, const uint8_t* bytes, size_t nbytes, void* const* planes,
						uint32_t numcomps, uint32_t stride, uint32_t plane_h, int reversible,
						uint32_t* status_out);

/* ---- quantisation parameters (host arithmetic only, no device) ------------------------------ */

/* The step sizes an encoder signals in QCD, one (expn, mant) per band in codestream order (LL,
 * then HL LH HH per resolution): param_qcd::generate, Part-1 branch
 * (src/lib/jp2/codestream/HTParams.cpp:194-252) and, for HT + 5/3, param_qcd::set_rev_quant
 * (:253-272, without the colour-transform bit, as the reference calls it).  expn/mant hold
 * 3*numres-2 entries. */
int grkb_qcd_generate(uint32_t numres, uint32_t prec, int reversible, int ht, uint32_t* expn,
					  uint32_t* mant);

/* Band step size and magnitude bit-plane count from one QCD entry:
 * Quantizer::setBandStepSizeAndBps (src/lib/jp2/codestream/Quantizer.cpp:26-66), including the
 * decoder's gain-0 rule for 9/7 and the HT lossy decoder's division by 2^(31-numbps). */
int grkb_band_quant(uint32_t expn, uint32_t mant, uint32_t orient, uint32_t prec, int reversible,
					int compress, int ht, uint32_t numgbits, uint32_t roishift, float* stepsize,
					uint32_t* numbps);

/* ---- encode side ---------------------------------------------------------------------------- */

/* One code block to encode: what CompressBlockExec carries (src/lib/jp2/t1/BlockExec.h:54-84,
 * filled by T1CompressScheduler::scheduleCompress, t1/T1CompressScheduler.cpp:31-154). */
typedef struct grkb_enc_cblk
{
	uint32_t x, y; /* top-left sample in the component's coefficient plane (Mallat layout) */
	uint16_t w, h;
	uint8_t comp;
	uint8_t orient; /* 0 LL, 1 HL, 2 LH, 3 HH */
	uint8_t level; /* numresolutions - 1 - resno: index into the band norms (T1.cpp:264-277) */
	uint8_t cblk_sty; /* GRKB_CBLKSTY_*; GRKB_CBLKSTY_HT selects the HT cleanup coder */
	uint8_t k_msbs; /* HT: band numbps (T1CompressScheduler.cpp:84 with cblk->numbps == 0) */
	uint8_t reserved[3];
	float stepsize; /* band step size: 9/7 quantiser divisor and distortion weight */
	uint32_t reserved2[2];
} grkb_enc_cblk; /* 32 bytes */

#define GRKB_ENC_MAX_PASSES 100 /* passes[] capacity of the reference (t1/Codeblock.h:144) */

/* One coding pass of a coded block: the fields of grk::CodePass the host's rate control and T2
 * read (rate = cumulative bytes, len = bytes of this pass, term = terminated,
 * distortion = cumulative weighted MSE decrease; T1.cpp:781-932). */
typedef struct grkb_pass
{
	double distortion;
	uint32_t rate;
	uint16_t len;
	uint8_t term;
	uint8_t reserved;
} grkb_pass; /* 16 bytes */

typedef struct grkb_coded
{
	uint64_t data_off; /* first byte of the block's stream in the output byte pool */
	uint32_t data_len;
	uint8_t numbps; /* MQ: floorlog2(max)+1-6; HT: 1 (T1HT.cpp:121-130) */
	uint8_t numpasses;
	uint16_t failed; /* 1: stream buffer overflow (the reference throws / asserts) */
	double distortion; /* total of the block (CompressBlockExec::distortion) */
} grkb_coded; /* 24 bytes */

/* Stage-level door for the block encoders (parity tests, and the call a maintainer binds in
 * place of T1CompressScheduler's worker loop): quantises (T1Part1::preCompress,
 * t1/t1_part1/T1Part1.cpp:37-88; T1HT::preCompress, t1/t1_ht/T1HT.cpp:56-108) and encodes
 * (T1::compress_cblk, t1/t1_part1/T1.cpp:781-932; ojph_encode_codeblock,
 * t1/t1_ht/coding/ojph_block_encoder.cpp:470-945) `nblocks` blocks of host coefficient planes.
 *   planes[c]     int32 (reversible) or float (irreversible) wavelet coefficients
 *   mct_norms     per-component weights of the distortion estimate (NULL: 1.0), T1.cpp:418-436
 *   rate_control  0: distortions are not computed (the reference skips them too)
 *   coded[i]      result of block i; passes[i * GRKB_ENC_MAX_PASSES + p] its passes
 *   bytes/cap     output byte pool (streams back to back in block order), *used = bytes written
 * HT with irreversible data is refused: the reference's own quantiser for that combination
 * reinterprets float bits as integers (T1HT.cpp:92-104). */
int grkb_t1_encode_host(const grkb_enc_cblk* blocks, uint32_t nblocks, const void* const* planes,
						uint32_t numcomps, uint32_t stride, uint32_t plane_h, int reversible,
						const double* mct_norms, uint32_t mct_numcomps, int rate_control,
						grkb_coded* coded, grkb_pass* passes, uint8_t* bytes, size_t cap,
						size_t* used);

/* Whole-tile encoder: the device side of TileProcessor::doCompress up to T2
 * (src/lib/jp2/tile/TileProcessor.cpp:202-232: dcLevelShiftCompress, mct_encode, dwt_encode,
 * t1_encode) for all components at once.  H2D of the int32 image planes, DC shift + forward MCT
 * fused into the first wavelet level, forward wavelet of every component, quantisation fused into
 * the block encoders, packing, D2H of the block table, the pass table and the coded bytes.
 * Outputs are as for grkb_t1_encode_host; rate allocation and T2 stay with the host. */
typedef struct grkb_encoder grkb_encoder;
int grkb_encoder_create(const grkb_tile_params* p, grkb_encoder** out);
void grkb_encoder_destroy(grkb_encoder* e);
int grkb_encoder_encode(grkb_encoder* e, const int32_t* const* image, uint32_t stride,
						const grkb_enc_cblk* blocks, uint32_t nblocks, const double* mct_norms,
						uint32_t mct_numcomps, int rate_control, grkb_coded* coded,
						grkb_pass* passes, uint8_t* bytes, size_t cap, size_t* used);
/* The same in two halves, so that several encoders (one per frame or tile, each with its own
 * stream) can be kept in flight and the copies of one overlap the kernels of another (the batch
 * path of grk_plugin_batch_compress): submit enqueues H2D, all kernels and the D2H of the block
 * table without waiting; wait blocks until they are done, then fetches the pass records that
 * exist (the first max(numpasses) of every row) and the coded bytes.  image, blocks, coded,
 * passes and bytes must stay valid until wait returns; page-locked memory (grkb_host_alloc)
 * makes the copies asynchronous. */
int grkb_encoder_submit(grkb_encoder* e, const int32_t* const* image, uint32_t stride,
						const grkb_enc_cblk* blocks, uint32_t nblocks, const double* mct_norms,
						uint32_t mct_numcomps, int rate_control, grkb_coded* coded,
						grkb_pass* passes, uint8_t* bytes, size_t cap);
int grkb_encoder_wait(grkb_encoder* e, size_t* used);
/* grkb_encoder_submit with the image in a narrow format (GRKB_SAMPLES_16 ..., see the decoder):
 * `image` holds grkb_samples_size(format, ...) bytes, widened on the device (sgnd: sign-extend). */
int grkb_encoder_submit_packed(grkb_encoder* e, const void* image, uint32_t format, int sgnd,
							   size_t row_bytes, const grkb_enc_cblk* blocks, uint32_t nblocks,
							   const double* mct_norms, uint32_t mct_numcomps, int rate_control,
							   grkb_coded* coded, grkb_pass* passes, uint8_t* bytes, size_t cap);
/* Re-runs the device stages of the last tile (image still resident), CUDA events on the
 * encoder's stream: ms[0] whole pass, ms[1] DC/MCT/wavelet, ms[2] block coding + packing. */
int grkb_encoder_time_device(grkb_encoder* e, uint32_t iters, uint32_t warmup, float ms[3]);

/* Page-locked host memory for byte pools and output planes: copies to and from it are true
 * asynchronous DMA (the decoder accepts any host pointer, pageable ones are staged by the
 * driver). */
void* grkb_host_alloc(size_t bytes);
void grkb_host_free(void* p);
/* Page-locks memory the caller already owns, so that results are written into it by DMA instead of
 * through the driver's staging buffer.  Only for page-aligned buffers that share no page with other
 * memory used in copies: the driver refuses a copy that straddles registered and unregistered
 * pages.  Failure is harmless (returns -1, the memory stays pageable). */
int grkb_host_register(void* p, size_t bytes);
int grkb_host_unregister(void* p);

/* C. diagnostics ------------------------------------------------------------------------------ */

/* Times the device-resident wavelet back end (forward == 0: inverse DWT [+ fused inverse MCT])
 * or front end (forward == 1) on synthetic planes already in HBM, with CUDA events on the
 * launching stream: `warmup` untimed then `iters` timed passes; *ms_per_iter = mean pass time.
 * Used by bench.py for the DWT roofline figure. */
int grkb_bench_dwt(uint32_t w, uint32_t h, uint32_t numres, int reversible, int forward,
				   uint32_t ncomp, int mct, uint32_t iters, uint32_t warmup, float* ms_per_iter);

/* Host <-> device link bandwidth with page-locked memory (what bounds every end-to-end figure,
 * SURVEY 8d): each direction alone and both at once, GB/s = 1e9 bytes/s. */
int grkb_bench_link(size_t bytes, uint32_t iters, double* h2d_gbs, double* d2h_gbs, double duplex_gbs[2]);

/* E. codestream level: T2 on our side ------------------------------------ */

/* Everything below parses JPEG 2000 codestreams itself (markers, tile parts, packet headers), so
 * that a whole image -- any number of tiles, layers, precincts, any progression order, POC, SOP /
 * EPH, several tile parts per tile -- goes from coded bytes to samples without the reference's T2.
 * It replaces, on the decode side, grk_decompress_read_header + grk_decompress
 * (src/lib/jp2/grok.h:1305-1345) up to the int32 planes of grk_image:
 * CodeStreamDecompress::readHeader / the marker readers (codestream/CodeStreamDecompress.cpp:
 * 1148-1860, codestream/markers/SIZMarker.cpp, SOTMarker.cpp), T2Decompress (t2/T2Decompress.cpp:
 * 28-566), PacketIter (t2/PacketIter.cpp:55-355), TagTree (t2/TagTree.h), BitIO (t2/BitIO.cpp) and the
 * tile compositing of CodeStreamDecompress::decompressTiles (:412-640).  Refused (error, nothing
 * decoded): sub-sampled components, more than GRKB_MAX_COMPS components, component-specific COC
 * geometry, PPM / PPT packed headers. */
typedef struct grkb_codestream grkb_codestream;

typedef struct grkb_codestream_info
{
	uint32_t x0, y0, x1, y1; /* image area on the reference grid (SIZ) */
	uint32_t tile_x0, tile_y0, tile_w, tile_h; /* tile grid origin and nominal tile size */
	uint32_t tiles_x, tiles_y;
	uint32_t numcomps;
	uint8_t prec[GRKB_MAX_COMPS], sgnd[GRKB_MAX_COMPS];
	/* main-header COD */
	uint32_t numres, numlayers, prog_order, reversible, mct, cblkw_exp, cblkh_exp, cblk_sty, csty;
	uint32_t ht; /* 1: HT block coder */
	uint32_t truncated; /* 1: the buffer ends inside a tile part */
	uint32_t rsiz;
	uint32_t tile_parts; /* tile-part bodies indexed */
	uint32_t features; /* GRKB_CS_* */
} grkb_codestream_info;
#define GRKB_CS_RGN 0x01 /* region of interest shift on some component */
#define GRKB_CS_QCC 0x02 /* component-specific quantisation */
#define GRKB_CS_COC 0x04 /* component-specific coding style (identical geometry: accepted) */
#define GRKB_CS_TILE_OVERRIDES 0x08 /* COD / QCD / ... in tile-part headers */
#define GRKB_CS_POC 0x10
#define GRKB_CS_MIXED_PRECISION 0x20 /* components differ in precision or signedness */

/* Reads the main header and indexes all tile parts of a raw codestream or a JP2 file (the jp2c box
 * is located).  `data` must stay valid and unchanged until grkb_codestream_close. */
int grkb_codestream_open(const uint8_t* data, size_t len, grkb_codestream** out);
void grkb_codestream_close(grkb_codestream* cs);
int grkb_codestream_get_info(const grkb_codestream* cs, grkb_codestream_info* info);

/* The descriptor tables of one tile after T2: what grkb_decoder_submit takes.  With zero_copy the
 * byte pool is a range of the caller's own buffer (every block's bytes are one run of the
 * codestream: any single-layer stream); otherwise the bytes were gathered into memory the
 * codestream object owns.  The tables stay valid until the tile is parsed again or the object is
 * closed. */
typedef struct grkb_tile_tables
{
	grkb_tile_params params; /* the tile the decoder sees: reduced rectangle, kept resolutions */
	const grkb_cblk* blocks;
	uint32_t nblocks;
	const grkb_seg* segs;
	uint32_t nsegs;
	const uint8_t* bytes;
	size_t nbytes;
	uint32_t zero_copy, packets, truncated;
} grkb_tile_tables;

/* T2 of one tile.  reduce: highest resolutions to leave out (cp_reduce of the reference,
 * tile/TileProcessor.cpp:288-317: the tile shrinks to ceil(rect / 2^reduce)); max_layers: quality
 * layers to keep, 0 = all.  Tiles may be parsed from different threads at once. */
int grkb_codestream_parse_tile(grkb_codestream* cs, uint32_t tileno, uint32_t reduce, uint32_t max_layers,
							   grkb_tile_tables* out);

/* Whole image: the device side of grk_decompress.  A decompressor owns `slots` tile decoders on
 * one device (decoders are cached per tile geometry), so the tiles of an image -- or, with one
 * slot per frame and several decompressors, several frames -- are in flight together. */
typedef struct grkb_decompressor grkb_decompressor;
int grkb_decompressor_create(int device, uint32_t slots, grkb_decompressor** out);
void grkb_decompressor_destroy(grkb_decompressor* d);

/* Parses `data` and enqueues every tile: H2D of its tables and bytes, T1, wavelet, MCT, D2H of its
 * rectangle into the output.  Returns once the last tile is enqueued; grkb_decompress_wait blocks
 * until all are done.  data, planes / out must stay valid until then (page-locked memory from
 * grkb_host_alloc makes the copies asynchronous).
 *   planes[c], stride   int32 planes of the (reduced) image, sized as the reference sizes its
 *                       grk_image: ceil((x1-x0)/2^r) x ceil((y1-y0)/2^r) (codestream/markers/
 *                       SIZMarker.cpp:54-57).  The tiles cover ceil(x1/2^r)-ceil(x0/2^r) of it; a
 *                       last row / column beyond that is left as the caller gave it (the reference
 *                       leaves zeros)
 * rejected_out: code blocks the block decoders refused (they decode as zeros). */
int grkb_decompress_submit(grkb_decompressor* d, const uint8_t* data, size_t len, uint32_t reduce,
						   uint32_t max_layers, int32_t* const* planes, uint32_t stride);
/* Windowed decode (grk_decompress_set_window, src/lib/jp2/grok.h:1321-1331 ->
 * CodeStreamDecompress::setDecompressWindow, codestream/CodeStreamDecompress.cpp:295-398, and the
 * partial wavelet of transform/WaveletReverse.cpp:1374-2230).  wx0..wy1: the window relative to the
 * image origin on the reference grid (0, 0, 0, 0: the whole image), clipped to the image.  Only the
 * tiles it touches are parsed and, inside them, only the code blocks whose samples reach the window
 * through the synthesis filters are decoded; planes receive the (ceil(x1/2^r) - ceil(x0/2^r)) x ...
 * samples of the window, identical to the same region of a full decode. */
int grkb_decompress_submit_window(grkb_decompressor* d, const uint8_t* data, size_t len, uint32_t reduce,
								  uint32_t max_layers, uint32_t wx0, uint32_t wy0, uint32_t wx1, uint32_t wy1,
								  int32_t* const* planes, uint32_t stride);
/* Tile sharding (tile t -> GPU t mod G, no collective): decodes tiles first_tile,
 * first_tile + tile_step, ... only; the other tiles' rectangles of `planes` are left alone.  Each
 * rank of a multi-GPU decode calls it with its own decompressor, (rank, world) and the same
 * codestream. */
int grkb_decompress_submit_tiles(grkb_decompressor* d, const uint8_t* data, size_t len, uint32_t reduce,
								 uint32_t max_layers, uint32_t first_tile, uint32_t tile_step, int32_t* const* planes,
								 uint32_t stride);
/* The same with the samples narrowed on the device (GRKB_SAMPLES_16 / _8, planar or interleaved;
 * opt-in): `out` holds grkb_samples_size(format, H, numcomps, row_bytes) bytes of the
 * whole (reduced) image. */
int grkb_decompress_submit_packed(grkb_decompressor* d, const uint8_t* data, size_t len, uint32_t reduce,
								  uint32_t max_layers, void* out, uint32_t format, size_t row_bytes);
int grkb_decompress_wait(grkb_decompressor* d, uint32_t* rejected_out);
/* submit + wait */
int grkb_decompress(grkb_decompressor* d, const uint8_t* data, size_t len, uint32_t reduce,
					uint32_t max_layers, int32_t* const* planes, uint32_t stride, uint32_t* rejected_out);
/* host milliseconds the last submit spent in marker + packet parsing (T2), and tiles enqueued */
int grkb_decompress_stats(const grkb_decompressor* d, double* t2_ms, uint32_t* tiles);

/* ---- the encode side of the codestream level ---------------------------------------------------
 * Image planes in, a complete JPEG 2000 / HTJ2K codestream out: the device tile encoder
 * (grkb_encoder_*) plus what the reference does around it -- main header (CodeStreamCompress::start and
 * the marker writers,
 * src/lib/jp2/codestream/CodeStreamCompress.cpp:640-1130), packet headers and bodies
 * (T2Compress::compressPacket / compressHeader, t2/T2Compress.cpp:112-330; TagTree::compress,
 * t2/TagTree.h:160-200), one tile part per tile.  It replaces grk_compress_init + grk_compress_start
 * + grk_compress + grk_compress_end (grok.h:1380-1420) for lossless J2K, HTJ2K lossless, all-passes
 * 9/7 streams, and one or several quality layers with rate targets (the reference's rate allocation,
 * see `rate` below), byte-identical to the stock encoder's output for the same parameters.  Not
 * written here: POC, tile-part division, TLM / PLT markers, the digital-cinema profiles' per-component
 * limits (those stay with the host behind grk_plugin_*). */
#define GRKB_MAX_LAYERS 100 /* maxCompressLayersGRK */
typedef struct grkb_compress_params
{
	uint32_t width, height, numcomps, prec, sgnd;
	uint32_t x0, y0; /* image origin on the reference grid */
	uint32_t tile_w, tile_h; /* 0: one tile */
	uint32_t tile_x0, tile_y0; /* tile grid origin (<= image origin) */
	uint32_t numres, cblkw_exp, cblkh_exp;
	uint32_t cblk_sty; /* GRKB_CBLKSTY_* (GRKB_CBLKSTY_HT alone for HTJ2K) */
	uint32_t reversible, mct;
	uint32_t prog_order; /* 0 LRCP, 1 RLCP, 2 RPCL, 3 PCRL, 4 CPRL */
	uint32_t csty; /* bit 1: SOP markers, bit 2: EPH markers */
	uint32_t numgbits; /* guard bits; 0: the reference's default (2, HT: 1) */
	uint8_t prec_w_exp[33], prec_h_exp[33]; /* log2 precinct size per resolution, lowest first (15: none) */
	char comment[64]; /* COM marker text; empty: "Created by grok_b200" */
	double rate; /* 0: keep every pass.  > 0: compression ratio of the single layer (-r of grk_compress):
				  * the block coders also deliver every pass's distortion and the passes that go into
				  * the layer are chosen by TileProcessor::pcrdBisectSimple's bisection of the slope
				  * threshold on the reference's own packet accounting (tile/
				  * TileProcessor.cpp:1195-1440, t2/T2Compress.cpp:58-110, :331-420), on the device
				  * behind the block coders (csrc/pcrd.cu; GRKB_PCRD=host: on the host) */
	uint32_t numlayers; /* 0 or 1: one layer at `rate`.  > 1: quality layers at layer_rates[] (-r a,b,c of
						 * grk_compress: compression ratios, decreasing; 0: every remaining pass), allocated
						 * layer by layer as pcrdBisectSimple does (each layer's search starts below the
						 * previous layer's threshold), on the host */
	double layer_rates[GRKB_MAX_LAYERS];
} grkb_compress_params;

typedef struct grkb_compressor grkb_compressor;
int grkb_compressor_create(int device, const grkb_compress_params* p, grkb_compressor** out);
void grkb_compressor_destroy(grkb_compressor* z);
/* planes[c]: int32 samples, `stride` per row.  out / cap: the codestream buffer; *used = bytes
 * written (or needed, with -2, when cap is too small). */
int grkb_compress(grkb_compressor* z, const int32_t* const* planes, uint32_t stride, uint8_t* out, size_t cap,
				  size_t* used);
/* milliseconds of the last call: device tile encodes (copies included), host T2 + markers */
int grkb_compress_stats(const grkb_compressor* z, double* device_ms, double* t2_ms);
/* rate allocation of the last call: tiles allocated on the device / on the host, bisection steps of the
 * last tile allocated on the device */
int grkb_compress_allocation(const grkb_compressor* z, uint32_t* tiles_on_device, uint32_t* tiles_on_host,
							 uint32_t* steps);

/* D. allocation-free door --------------------------------------------------------------------- */

/* No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * allocation).  Multi-level inverse wavelet of ONE tile-component on DEVICE buffers, the same
 * kernels and results as grkb_dwt_host(forward = 0) / grkb_decoder_*, with the two ping-pong planes
 * of the intermediate resolutions taken from a file-scope __device__ array of
 * GRKB_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit.  The call
 * performs no cudaMalloc / cudaFree / allocating library routine; it enqueues on `stream`
 * (a cudaStream_t; NULL = the default stream) and returns without synchronising.
 *   coeff   device: dequantised coefficients, tile buffer layout, coeff_stride samples per row
 *           (read only)
 *   image   device: receives the (x1-x0) x (y1-y0) samples, image_stride samples per row; must not
 *           alias coeff
 * Limit: 2 x stride32(ceil(w/2)) x ceil(h/2) x 4 bytes <= GRKB_NOALLOC_SCRATCH_BYTES, i.e.
 * tile-components up to 2048 x 2048; larger ones return -2 with a message in grkb_last_error().
 * All calls on one GPU share the array: calls that overlap on the GPU (different streams, or
 * different threads without ordering) race -- serialise them.  Nothing in the array has to start
 * at zero (every level overwrites what the next one reads), so nothing is cleared between calls. */
#define GRKB_NOALLOC_SCRATCH_BYTES (8u << 20)
int grkb_idwt_device_noalloc(const void* coeff, uint32_t coeff_stride, void* image,
							 uint32_t image_stride, uint32_t x0, uint32_t y0, uint32_t x1,
							 uint32_t y1, uint32_t numres, int reversible, void* stream);

#pragma GCC visibility pop
#ifdef __cplusplus
}
#endif
#endif
