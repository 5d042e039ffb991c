// t1.cuh -- shared declarations of the block-coder kernels (EBCOT/MQ and HTJ2K), sm_100a.
//
// Stripe-column state word used by the MQ coder kernels (one 32-bit word per column per 4-row
// stripe, own layout; the reference's is in t1/t1_part1/T1.cpp:27-131):
//   bits  0..17  significance of the 3x6 neighbourhood, row-major: row j (j = -1..4 relative to
//                the stripe) occupies bits 3(j+1)..3(j+1)+2 as [west, centre, east]
//   bits 18..23  sign (1 = negative) of the centre column, rows -1..4
//   bits 24..27  "visited in this bit-plane" (pi) of rows 0..3
//   bits 28..31  "already refined" (mu) of rows 0..3
// so that (word >> 3r) & 0x1FF is the 3x3 neighbourhood of row r, which indexes the
// zero-coding table directly.
#pragma once
#include "common.cuh"
#include "../../include/grk_b200.h"

#define T1_SIG_C(r) (1u << (3 * ((r) + 1) + 1))
#define T1_NEIGH 0x1EFu
#define T1_SGN(r) (1u << (18 + (r) + 1))
#define T1_PI(r) (1u << (24 + (r)))
#define T1_MU(r) (1u << (28 + (r)))
#define T1_PI_ALL (0xFu << 24)
#define T1_SIGC_ALL (T1_SIG_C(0) | T1_SIG_C(1) | T1_SIG_C(2) | T1_SIG_C(3))

enum
{
	T1_CTX_ZC = 0,
	T1_CTX_SC = 9,
	T1_CTX_MAG = 14,
	T1_CTX_AGG = 17,
	T1_CTX_UNI = 18,
	T1_NCTX = 19
};

// Tables derived on the host from the rules of ISO/IEC 15444-1 Annex C/D at library start.
struct T1Tables
{
	uint32_t mq[94]; // state 2*i+mps: qe | next-if-mps << 16 | next-if-lps << 24 (table C.2)
	uint8_t zc[4 * 512]; // [orientation][3x3 neighbourhood] -> context 0..8 (table D.1)
	uint8_t sc[256]; // [sigW sigE sigN sigS sgnW sgnE sgnN sgnS] -> context | xorbit << 7 (D.3)
	int16_t nmsedec[4][128]; // distortion tables: sig, sig0, ref, ref0 (encoder)
};

const T1Tables* grkb_t1_tables_device(); // device copy (per device), nullptr on failure
const T1Tables& grkb_t1_tables_host();

// launchers
struct T1Launch
{
	const grkb_cblk* blocks; // device
	const grkb_seg* segs; // device
	const uint8_t* bytes; // device
	const uint32_t* order; // device: block indices, longest first (may be null)
	uint32_t nblocks;
	void* planes[GRKB_MAX_COMPS]; // device coefficient planes
	uint32_t stride;
	int reversible;
	uint32_t* counter; // device work counter (zeroed by the launcher)
	uint32_t* status; // device: count of rejected blocks
	uint32_t max_area; // max w*h over MQ blocks
	uint32_t max_flag_words; // max (w+2)*(ceil(h/4)+2) over MQ blocks
	uint32_t max_ht_w; // max width over HT blocks
	uint32_t ht_refine; // some HT block has more than one codeword segment (SigProp / MagRef)
	uint32_t concurrent; // other tile decoders of this process with block decoding in flight on the device
};
// encode side: every block codes into its own slot of `slots`; a second kernel packs the
// streams back to back (block order) into `packed` and fills coded[].data_off
struct T1EncLaunch
{
	const grkb_enc_cblk* blocks; // device
	uint32_t nblocks;
	const void* planes[GRKB_MAX_COMPS]; // device coefficient planes
	uint32_t stride;
	int reversible;
	double mct_norms[GRKB_MAX_COMPS];
	uint32_t mct_numcomps;
	int rate_control;
	grkb_coded* coded; // device, nblocks
	grkb_pass* passes; // device, nblocks * GRKB_ENC_MAX_PASSES
	uint8_t* slots; // device, nblocks * slot_bytes (each slot starts with 16 zero bytes)
	uint32_t slot_bytes;
	uint8_t* packed; // device
	uint64_t packed_cap;
	uint64_t* packed_used; // device: total bytes, or ~0 if the pool was too small
	uint32_t* counter; // device work counter (zeroed by the launcher)
	uint32_t max_area; // max w*h over the blocks
	uint32_t max_flag_words; // max (w+2)*(ceil(h/4)+2) over MQ blocks
	uint32_t max_w; // widest block
	// MQ blocks are coded in two kernels: context modelling leaves every block's (context,
	// decision) symbols in chunks of this pool, the arithmetic coder reads them back
	uint8_t* sym_pool; // device
	uint64_t sym_cap; // bytes
	uint32_t* sym_ctrl; // device: [0] chunks handed out, [1] set when the pool ran out
	uint64_t total_area; // sum of w*h over the MQ blocks (sizes the pool)
	// the arithmetic coder takes the blocks longest first: symbols per block (left by the context
	// modelling), the block indices in that order, and [0] how many of the longest blocks get a
	// warp of their own
	uint32_t* sym_total; // device, nblocks
	uint32_t* code_order; // device, nblocks
	uint32_t* code_ctrl; // device
};
int grkb_launch_t1_mq_encode(const T1EncLaunch& l, cudaStream_t s);
int grkb_launch_t1_ht_encode(const T1EncLaunch& l, cudaStream_t s);
int grkb_launch_t1_pack(const T1EncLaunch& l, cudaStream_t s);
int grkb_launch_t1_mq_decode(const T1Launch& l, cudaStream_t s);
int grkb_launch_t1_ht_decode(const T1Launch& l, cudaStream_t s);
