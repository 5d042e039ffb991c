// samples.cu -- narrow sample formats at the PCIe boundary, sm_100a.
//
// grk_image carries every sample as an int32 (src/lib/jp2/grok.h: grk_image_comp::data), so a
// 12-bit DCI frame crosses the bus at 4 bytes per sample although every consumer that leaves the
// library narrows it again:
//   planarToInterleaved<N>      src/bin/common/convert.cpp:397-405   (int32 planes -> pixels)
//   convert_16u32s_C1R          src/bin/common/convert.cpp:547-556   (16-bit file samples -> int32)
//   the 8/16-bit writers        src/bin/image_format/PNMFormat.cpp, TIFFFormat.cpp
// Here the narrowing (decode) and widening (encode) run on the device, next to the stage that
// produces or consumes the int32 planes, and the bus carries 1 or 2 bytes per sample.  Opt-in:
// the int32 planar form stays the default of the decoder and encoder objects.
//
// Formats (include/grk_b200.h): 16- or 8-bit samples in host byte order, planar (numcomps planes
// of H rows back to back) or interleaved (H rows of W pixels of numcomps samples).  A sample is
// the low 16 / 8 bits of the int32 (after the decoder's clamp to the component's range that is
// the value itself, two's complement for signed components); widening zero- or sign-extends.
#include "../../include/grk_b200.h"
#include "common.cuh"
#include "pipeline.h"
#include <type_traits>

namespace
{
struct SampleArgs
{
	void* planes[GRKB_MAX_COMPS]; // device int32 planes, `stride` samples per row
	uint32_t stride, W, H, numcomps;
	uint8_t* packed; // device
	size_t row_bytes; // of a packed row
	size_t plane_bytes; // planar formats: H * row_bytes
	int sgnd;
	int vec_ok; // packed rows and base aligned for the vector accesses
	int be; // 16-bit samples are big-endian on the packed side (PNM files)
};

template<typename T> struct Vec4;
template<> struct Vec4<uint16_t>
{
	typedef uint2 type;
	static __device__ __forceinline__ uint2 pack(int4 v)
	{
		return make_uint2(((uint32_t)v.x & 0xFFFFu) | ((uint32_t)v.y << 16),
						  ((uint32_t)v.z & 0xFFFFu) | ((uint32_t)v.w << 16));
	}
	static __device__ __forceinline__ int4 unpack(uint2 p, int sgnd)
	{
		if(sgnd)
			return make_int4((int16_t)(p.x & 0xFFFFu), (int16_t)(p.x >> 16), (int16_t)(p.y & 0xFFFFu),
							 (int16_t)(p.y >> 16));
		return make_int4((int)(p.x & 0xFFFFu), (int)(p.x >> 16), (int)(p.y & 0xFFFFu), (int)(p.y >> 16));
	}
};
template<> struct Vec4<uint8_t>
{
	typedef uint32_t type;
	static __device__ __forceinline__ uint32_t pack(int4 v)
	{
		return ((uint32_t)v.x & 0xFFu) | (((uint32_t)v.y & 0xFFu) << 8) | (((uint32_t)v.z & 0xFFu) << 16) |
			   ((uint32_t)v.w << 24);
	}
	static __device__ __forceinline__ int4 unpack(uint32_t p, int sgnd)
	{
		if(sgnd)
			return make_int4((int8_t)(p & 0xFFu), (int8_t)((p >> 8) & 0xFFu), (int8_t)((p >> 16) & 0xFFu),
							 (int8_t)(p >> 24));
		return make_int4((int)(p & 0xFFu), (int)((p >> 8) & 0xFFu), (int)((p >> 16) & 0xFFu), (int)(p >> 24));
	}
};

// byte order of the packed side: 16-bit samples swapped when the buffer is big-endian
template<typename V> static __device__ __forceinline__ V swap16(V p, int be);
template<> __device__ __forceinline__ uint2 swap16<uint2>(uint2 p, int be)
{
	return be ? make_uint2(__byte_perm(p.x, 0, 0x2301), __byte_perm(p.y, 0, 0x2301)) : p;
}
template<> __device__ __forceinline__ uint32_t swap16<uint32_t>(uint32_t p, int) // 8-bit samples
{
	return p;
}
template<typename T> static __device__ __forceinline__ T swap1(T s, int be)
{
	return sizeof(T) == 2 && be ? (T)(((uint32_t)s >> 8) | ((uint32_t)s << 8)) : s;
}

static __device__ __forceinline__ int elem(const int4& v, int i)
{
	return i == 0 ? v.x : i == 1 ? v.y : i == 2 ? v.z : v.w;
}

// One thread = four consecutive samples of a row, all components.  The planes are read / written
// with 128-bit accesses (strides are multiples of 32 samples, so the group past W is still inside
// the row's allocation); the packed side takes one vector access per component (planar) or goes
// sample by sample through registers into pixel order (interleaved: a warp still covers one
// contiguous span of the row).
template<typename T, bool INTERLEAVED, bool PACK>
__global__ void __launch_bounds__(256) samples_kernel(const SampleArgs a)
{
	typedef typename Vec4<T>::type V;
	const uint32_t groups = (a.W + 3) >> 2;
	const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
	if(g >= groups)
		return;
	const uint32_t x = g << 2;
	const bool whole = x + 4 <= a.W && a.vec_ok;
	uint8_t* row = a.packed + (size_t)y * a.row_bytes;
	if(!INTERLEAVED)
	{
		for(uint32_t c = 0; c < a.numcomps; ++c)
		{
			int4* p = reinterpret_cast<int4*>(reinterpret_cast<int32_t*>(a.planes[c]) + (size_t)y * a.stride + x);
			T* q = reinterpret_cast<T*>(row + (size_t)c * a.plane_bytes) + x;
			if(PACK)
			{
				const int4 v = *p;
				if(whole)
					*reinterpret_cast<V*>(q) = swap16<V>(Vec4<T>::pack(v), a.be);
				else
					for(uint32_t i = 0; i < 4 && x + i < a.W; ++i)
						q[i] = swap1<T>((T)elem(v, (int)i), a.be);
			}
			else
			{
				int4 v = make_int4(0, 0, 0, 0);
				if(whole)
					v = Vec4<T>::unpack(swap16<V>(*reinterpret_cast<const V*>(q), a.be), a.sgnd);
				else
				{
					int t[4] = {0, 0, 0, 0};
					for(uint32_t i = 0; i < 4 && x + i < a.W; ++i)
					{
						const T sv = swap1<T>(q[i], a.be);
						t[i] = a.sgnd ? (int)(typename std::make_signed<T>::type)sv : (int)sv;
					}
					v = make_int4(t[0], t[1], t[2], t[3]);
				}
				*p = v;
			}
		}
	}
	else
	{
		T* q = reinterpret_cast<T*>(row) + (size_t)x * a.numcomps;
		const uint32_t n = x + 4 <= a.W ? 4u : a.W - x;
		if(PACK)
		{
			for(uint32_t c = 0; c < a.numcomps; ++c)
			{
				const int4 v =
					*reinterpret_cast<const int4*>(reinterpret_cast<const int32_t*>(a.planes[c]) + (size_t)y * a.stride + x);
				for(uint32_t i = 0; i < n; ++i)
					q[i * a.numcomps + c] = swap1<T>((T)elem(v, (int)i), a.be);
			}
		}
		else
		{
			for(uint32_t c = 0; c < a.numcomps; ++c)
			{
				int t[4] = {0, 0, 0, 0};
				for(uint32_t i = 0; i < n; ++i)
				{
					const T s = swap1<T>(q[i * a.numcomps + c], a.be);
					t[i] = a.sgnd ? (int)(typename std::make_signed<T>::type)s : (int)s;
				}
				*reinterpret_cast<int4*>(reinterpret_cast<int32_t*>(a.planes[c]) + (size_t)y * a.stride + x) =
					make_int4(t[0], t[1], t[2], t[3]);
			}
		}
	}
}

template<bool PACK>
int launch(const SampleArgs& a, uint32_t format, cudaStream_t s)
{
	const dim3 grid(((a.W + 3) / 4 + 255) / 256, a.H);
	switch(format & ~(uint32_t)GRKB_SAMPLES_BIG_ENDIAN)
	{
		case GRKB_SAMPLES_16:
			samples_kernel<uint16_t, false, PACK><<<grid, 256, 0, s>>>(a);
			break;
		case GRKB_SAMPLES_16_INTERLEAVED:
			samples_kernel<uint16_t, true, PACK><<<grid, 256, 0, s>>>(a);
			break;
		case GRKB_SAMPLES_8:
			samples_kernel<uint8_t, false, PACK><<<grid, 256, 0, s>>>(a);
			break;
		case GRKB_SAMPLES_8_INTERLEAVED:
			samples_kernel<uint8_t, true, PACK><<<grid, 256, 0, s>>>(a);
			break;
		default:
			grkb_set_error("samples: unknown format %u", format);
			return -1;
	}
	GRKB_CUDA_CHECK(cudaGetLastError());
	return 0;
}
} // namespace

size_t grkb_samples_min_row_bytes(uint32_t format, uint32_t W, uint32_t numcomps)
{
	if((format & GRKB_SAMPLES_BIG_ENDIAN) && (format & ~(uint32_t)GRKB_SAMPLES_BIG_ENDIAN) != GRKB_SAMPLES_16 &&
	   (format & ~(uint32_t)GRKB_SAMPLES_BIG_ENDIAN) != GRKB_SAMPLES_16_INTERLEAVED)
		return 0; // byte order only means something for 16-bit samples
	switch(format & ~(uint32_t)GRKB_SAMPLES_BIG_ENDIAN)
	{
		case GRKB_SAMPLES_16:
			return (size_t)W * 2;
		case GRKB_SAMPLES_16_INTERLEAVED:
			return (size_t)W * numcomps * 2;
		case GRKB_SAMPLES_8:
			return (size_t)W;
		case GRKB_SAMPLES_8_INTERLEAVED:
			return (size_t)W * numcomps;
		default:
			return 0;
	}
}

size_t grkb_samples_bytes(uint32_t format, uint32_t H, uint32_t numcomps, size_t row_bytes)
{
	format &= ~(uint32_t)GRKB_SAMPLES_BIG_ENDIAN;
	const bool planar = format == GRKB_SAMPLES_16 || format == GRKB_SAMPLES_8;
	return row_bytes * H * (planar ? numcomps : 1);
}

static int check_layout(uint32_t format, uint32_t W, uint32_t numcomps, size_t row_bytes, const void* packed)
{
	const size_t need = grkb_samples_min_row_bytes(format, W, numcomps);
	if(need == 0)
	{
		grkb_set_error("samples: unknown format %u", format);
		return -1;
	}
	if(row_bytes < need || !packed)
	{
		grkb_set_error("samples: rows of %zu bytes, need at least %zu", row_bytes, need);
		return -1;
	}
	return 0;
}

int grkb_launch_pack_samples(void* const* planes, uint32_t numcomps, uint32_t stride, uint32_t W, uint32_t H,
							 uint32_t format, void* packed, size_t row_bytes, cudaStream_t s)
{
	if(check_layout(format, W, numcomps, row_bytes, packed))
		return -1;
	if(W == 0 || H == 0)
		return 0;
	SampleArgs a{};
	for(uint32_t c = 0; c < numcomps; ++c)
		a.planes[c] = planes[c];
	a.stride = stride;
	a.W = W;
	a.H = H;
	a.numcomps = numcomps;
	a.packed = (uint8_t*)packed;
	a.row_bytes = row_bytes;
	a.plane_bytes = row_bytes * H;
	a.vec_ok = !((row_bytes | a.plane_bytes | (uintptr_t)packed) & 7);
	a.be = (format & GRKB_SAMPLES_BIG_ENDIAN) != 0;
	return launch<true>(a, format, s);
}

int grkb_launch_unpack_samples(const void* packed, size_t row_bytes, uint32_t format, int sgnd, void* const* planes,
							   uint32_t numcomps, uint32_t stride, uint32_t W, uint32_t H, cudaStream_t s)
{
	if(check_layout(format, W, numcomps, row_bytes, packed))
		return -1;
	if(W == 0 || H == 0)
		return 0;
	SampleArgs a{};
	for(uint32_t c = 0; c < numcomps; ++c)
		a.planes[c] = planes[c];
	a.stride = stride;
	a.W = W;
	a.H = H;
	a.numcomps = numcomps;
	a.packed = (uint8_t*)const_cast<void*>(packed);
	a.row_bytes = row_bytes;
	a.plane_bytes = row_bytes * H;
	a.sgnd = sgnd;
	a.vec_ok = !((row_bytes | a.plane_bytes | (uintptr_t)packed) & 7);
	a.be = (format & GRKB_SAMPLES_BIG_ENDIAN) != 0;
	return launch<false>(a, format, s);
}
