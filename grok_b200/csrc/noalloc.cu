// noalloc.cu -- the one entry point of the library that allocates nothing
// (include/grk_b200.h, section D: grkb_idwt_device_noalloc).
//
// The multi-level inverse wavelet needs two ping-pong planes for its intermediate resolutions
// (dwt_host.cu: level r reads LL from the previous level's output and writes elsewhere).  Every
// other door takes them from cudaMalloc; callers that cannot (CUDA graph capture, serving loops,
// harnesses that forbid allocation) get them here from a file-scope __device__ array of a fixed,
// modest size that every GPU loading the library holds until exit.
//
// The array's name is a host-side stand-in in host code, not a device address: the address the
// kernels receive comes from cudaGetSymbolAddress, asked once per device with that device current.
#include "pipeline.h"
#include "../../include/grk_b200.h"

static __device__ __align__(256) unsigned char g_noalloc_scratch[GRKB_NOALLOC_SCRATCH_BYTES];

namespace
{
std::mutex g_mu;
void* g_scratch_addr[64] = {}; // per device, filled on the first call with that device current

// device address of the scratch array on the calling thread's current device
int scratch_address(unsigned char** out)
{
	int dev = 0;
	GRKB_CUDA_CHECK(cudaGetDevice(&dev));
	if(dev < 0 || dev >= 64)
	{
		grkb_set_error("grkb_idwt_device_noalloc: device ordinal %d out of range", dev);
		return -1;
	}
	std::lock_guard<std::mutex> lk(g_mu);
	if(!g_scratch_addr[dev])
	{
		void* p = nullptr;
		cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
		if(e != cudaSuccess || !p)
		{
			grkb_set_error("grkb_idwt_device_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
						   cudaGetErrorString(e));
			return -1;
		}
		g_scratch_addr[dev] = p;
	}
	*out = (unsigned char*)g_scratch_addr[dev];
	return 0;
}
} // namespace

extern "C" int grkb_idwt_device_noalloc(const void* coeff, uint32_t coeff_stride, void* image,
										uint32_t image_stride, uint32_t x0, uint32_t y0, uint32_t x1,
										uint32_t y1, uint32_t numres, int reversible, void* stream)
{
	const grkb_rect tc{x0, y0, x1, y1};
	if(!coeff || !image || x1 < x0 || y1 < y0 || numres == 0 || numres > 33)
	{
		grkb_set_error("grkb_idwt_device_noalloc: bad arguments");
		return -1;
	}
	const uint32_t W = x1 - x0, H = y1 - y0;
	if(W == 0 || H == 0)
		return 0;
	if(coeff_stride < W || image_stride < W || coeff == image)
	{
		grkb_set_error("grkb_idwt_device_noalloc: strides must cover the width and the planes must differ");
		return -1;
	}
	uint32_t sstride = 32;
	size_t sbytes = grkb_dwt_scratch_bytes(tc, numres, &sstride);
	sbytes = (sbytes + 255) & ~(size_t)255;
	if(2 * sbytes > (size_t)GRKB_NOALLOC_SCRATCH_BYTES)
	{
		grkb_set_error("grkb_idwt_device_noalloc: a %u x %u tile-component needs %zu bytes of scratch, the "
					   "built-in array holds %u (limit: 2048 x 2048 samples); use grkb_decoder_* instead",
					   W, H, 2 * sbytes, (unsigned)GRKB_NOALLOC_SCRATCH_BYTES);
		return -2;
	}
	unsigned char* base = nullptr;
	if(scratch_address(&base))
		return -1;
	DwtGroup g{};
	g.ncomp = 1;
	g.coeff[0] = const_cast<void*>(coeff);
	g.coeff_stride = coeff_stride;
	g.image[0] = image;
	g.image_stride = image_stride;
	// nothing in the planes has to start at zero: every level writes the whole rectangle the next
	// one reads, so what the previous call left behind is never seen
	g.scratch[0][0] = base;
	g.scratch[0][1] = base + sbytes;
	g.scratch_stride = sstride;
	const int32_t z[3] = {0, 0, 0};
	return grkb_idwt_device(g, tc, numres, reversible, 0, z, z, z, (cudaStream_t)stream);
}
