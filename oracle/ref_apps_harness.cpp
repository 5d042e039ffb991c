/*
 * ref_apps_harness.cpp -- extern "C" doors into the UNMODIFIED sample conversion loops of the
 * reference's applications (src/bin/common/convert.cpp), compiled where they lie by oracle/Makefile
 * into oracle/_ref/libgrok_ref_apps.so.  TEST INFRASTRUCTURE ONLY: it holds no arithmetic of its
 * own, it only calls the reference's functions so that oracle/samples.c can be pinned to them.
 */
#include <stddef.h>
#include <stdint.h>
#include "grok.h"
#include "convert.h"

extern "C" {

/* planarToInterleaved<N> (convert.cpp:397-405) through the reference's own dispatch table */
int ref_planar_to_interleaved(const int32_t* const* src, uint32_t numcomps, int32_t* dst, size_t length,
							  int32_t adjust)
{
	if(numcomps == 0 || numcomps > 9)
		return -1;
	cvtPlanarToInterleaved_LUT[numcomps](src, dst, length, adjust);
	return 0;
}
/* interleavedToPlanar<N> (convert.cpp:360-384) */
int ref_interleaved_to_planar(const int32_t* src, int32_t* const* dst, uint32_t numcomps, size_t length)
{
	if(numcomps == 0 || numcomps > 9)
		return -1;
	cvtInterleavedToPlanar_LUT[numcomps](src, dst, length);
	return 0;
}
/* convert_tif_32sto16u (convert.cpp:1158-1163) */
void ref_32s_to_16u(const int32_t* src, uint8_t* dst, size_t length)
{
	convert_tif_32sto16u(src, dst, length);
}
/* cvtFrom32_LUT[8] = convert_32s8u_C1R (convert.cpp:564 on) */
void ref_32s_to_8u(const int32_t* src, uint8_t* dst, size_t length)
{
	cvtFrom32_LUT[8](src, dst, length);
}
/* convert_16u32s_C1R (convert.cpp:547-556): big-endian file bytes -> int32 */
void ref_16u_to_32s(const uint8_t* src, int32_t* dst, size_t length)
{
	convert_16u32s_C1R(src, dst, length, false);
}
/* cvtTo32_LUT[8] = convert_8u32s_C1R */
void ref_8u_to_32s(const uint8_t* src, int32_t* dst, size_t length)
{
	cvtTo32_LUT[8](src, dst, length, false);
}
}
