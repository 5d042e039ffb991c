// dwt_host.cu -- level sequencing for the wavelet kernels and the stage-level C-ABI on host
// buffers (include/grk_b200.h, section A).
//
// Device data layout: per tile-component one coefficient plane in the reference's Mallat
// layout, one full-resolution image plane, and two quarter-size scratch planes.  The inverse
// transform reads LL from the previous level's output and HL/LH/HH from the coefficient plane
// and writes to the other scratch plane (the last level writes the image plane), so no level
// ever reads what another CTA of the same launch writes.  The forward transform is the mirror.
#include "pipeline.h"
#include "../../include/grk_b200.h"
#include <ctype.h>
#include <sched.h>
#include <stdarg.h>
#include <vector>
#include <string.h>

static thread_local char g_err[512] = "";

void grkb_set_error(const char* fmt, ...)
{
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(g_err, sizeof(g_err), fmt, ap);
	va_end(ap);
}

size_t grkb_dwt_scratch_bytes(grkb_rect tc, uint32_t numres, uint32_t* stride_out)
{
	if(numres < 3)
	{
		if(stride_out)
			*stride_out = 32;
		return 0;
	}
	grkb_rect r = grkb_band_rect(tc, 1, 0); // resolution numres-2
	uint32_t w = r.x1 - r.x0, h = r.y1 - r.y0;
	uint32_t s = grkb_stride(w ? w : 1);
	if(stride_out)
		*stride_out = s;
	return (size_t)s * (h ? h : 1) * sizeof(int32_t);
}

int grkb_idwt_device(const DwtGroup& g, grkb_rect tc, uint32_t numres, int reversible, int fin,
					 const int32_t* shift, const int32_t* lo, const int32_t* hi, cudaStream_t s,
					 cudaEvent_t* ev_last, uint32_t* launches)
{
	const uint32_t W = tc.x1 - tc.x0, Hh = tc.y1 - tc.y0;
	if(W == 0 || Hh == 0)
		return 0;
	if(numres <= 1)
	{
		// no wavelet level: coefficients are the samples; apply the finishing stage on a copy
		for(int c = 0; c < g.ncomp; ++c)
			GRKB_CUDA_CHECK(cudaMemcpy2DAsync(g.image[c], (size_t)g.image_stride * 4, g.coeff[c],
											  (size_t)g.coeff_stride * 4, (size_t)W * 4, Hh,
											  cudaMemcpyDeviceToDevice, s));
		if(fin == 0)
			return 0;
		MctArgs m{};
		m.n = (uint64_t)g.image_stride * Hh;
		for(int c = 0; c < 3; ++c)
		{
			m.shift[c] = shift[c < g.ncomp ? c : 0];
			m.lo[c] = lo[c < g.ncomp ? c : 0];
			m.hi[c] = hi[c < g.ncomp ? c : 0];
		}
		if(fin == 2)
		{
			for(int c = 0; c < 3; ++c)
				m.c[c] = g.image[c];
			return grkb_launch_mct(m, reversible ? 0 : 1, s);
		}
		for(int c = 0; c < g.ncomp; ++c)
		{
			MctArgs d = m;
			d.c[0] = g.image[c];
			d.shift[0] = shift[c];
			d.lo[0] = lo[c];
			d.hi[0] = hi[c];
			if(grkb_launch_dc(d, reversible ? 0 : 1, s))
				return -1;
		}
		return 0;
	}
	const void* ll[3];
	uint32_t ll_stride = g.coeff_stride;
	for(int c = 0; c < g.ncomp; ++c)
		ll[c] = g.coeff[c];
	grkb_rect prev = grkb_band_rect(tc, numres - 1, 0);
	int ping = 0;
	for(uint32_t res = 1; res < numres; ++res)
	{
		grkb_rect cur = grkb_band_rect(tc, numres - 1 - res, 0);
		DwtLevelArgs a{};
		a.rw = cur.x1 - cur.x0;
		a.rh = cur.y1 - cur.y0;
		a.sw = prev.x1 - prev.x0;
		a.sh = prev.y1 - prev.y0;
		a.px = cur.x0 & 1;
		a.py = cur.y0 & 1;
		prev = cur;
		const bool last = (res == numres - 1);
		if(a.rw == 0 || a.rh == 0)
			continue; // WaveletReverse.cpp:833-834
		for(int c = 0; c < g.ncomp; ++c)
		{
			a.pl[c].ll = ll[c];
			a.pl[c].ll_stride = ll_stride;
			a.pl[c].bands = g.coeff[c];
			a.pl[c].b_stride = g.coeff_stride;
			a.pl[c].full = last ? g.image[c] : g.scratch[c][ping];
			a.pl[c].f_stride = last ? g.image_stride : g.scratch_stride;
		}
		int lfin = 0;
		if(last)
		{
			lfin = fin;
			for(int c = 0; c < 3; ++c)
			{
				a.shift[c] = shift[c < g.ncomp ? c : 0];
				a.lo[c] = lo[c < g.ncomp ? c : 0];
				a.hi[c] = hi[c < g.ncomp ? c : 0];
			}
		}
		if(last && ev_last)
			cudaEventRecord(ev_last[0], s);
		if(launches)
			*launches += (g.ncomp == 3) ? 1 : (uint32_t)g.ncomp;
		if(g.ncomp == 3)
		{
			if(grkb_launch_idwt_level(a, reversible, 3, lfin == 2 ? 2 : 0, s))
				return -1;
			if(lfin == 1)
			{
				grkb_set_error("idwt: per-component DC shift on a 3-component group is done by the caller");
				return -1;
			}
		}
		else
		{
			for(int c = 0; c < g.ncomp; ++c)
			{
				DwtLevelArgs b = a;
				b.pl[0] = a.pl[c];
				b.shift[0] = a.shift[c];
				b.lo[0] = a.lo[c];
				b.hi[0] = a.hi[c];
				if(grkb_launch_idwt_level(b, reversible, 1, lfin == 1 ? 1 : 0, s))
					return -1;
			}
		}
		if(last && ev_last)
			cudaEventRecord(ev_last[1], s);
		for(int c = 0; c < g.ncomp; ++c)
			ll[c] = a.pl[c].full;
		ll_stride = last ? g.image_stride : g.scratch_stride;
		ping ^= 1;
	}
	return 0;
}

int grkb_fdwt_device(const DwtGroup& g, grkb_rect tc, uint32_t numres, int reversible, int fin,
					 const int32_t* shift, cudaStream_t s)
{
	const uint32_t W = tc.x1 - tc.x0, Hh = tc.y1 - tc.y0;
	if(W == 0 || Hh == 0)
		return 0;
	if(numres <= 1)
	{
		for(int c = 0; c < g.ncomp; ++c)
			GRKB_CUDA_CHECK(cudaMemcpy2DAsync(g.coeff[c], (size_t)g.coeff_stride * 4, g.image[c],
											  (size_t)g.image_stride * 4, (size_t)W * 4, Hh,
											  cudaMemcpyDeviceToDevice, s));
		if(fin == 0)
			return 0;
		MctArgs m{};
		m.n = (uint64_t)g.coeff_stride * Hh;
		for(int c = 0; c < 3; ++c)
			m.shift[c] = shift[c < g.ncomp ? c : 0];
		if(fin == 2)
		{
			for(int c = 0; c < 3; ++c)
				m.c[c] = g.coeff[c];
			return grkb_launch_mct(m, reversible ? 2 : 3, s);
		}
		for(int c = 0; c < g.ncomp; ++c)
		{
			MctArgs d = m;
			d.c[0] = g.coeff[c];
			d.shift[0] = shift[c];
			if(grkb_launch_dc(d, reversible ? 2 : 3, s))
				return -1;
		}
		return 0;
	}
	const void* src[3];
	uint32_t src_stride = g.image_stride;
	for(int c = 0; c < g.ncomp; ++c)
		src[c] = g.image[c];
	int ping = 0;
	bool first = true;
	for(uint32_t res = numres - 1; res >= 1; --res)
	{
		grkb_rect cur = grkb_band_rect(tc, numres - 1 - res, 0);
		grkb_rect low = grkb_band_rect(tc, numres - res, 0);
		DwtLevelArgs a{};
		a.rw = cur.x1 - cur.x0;
		a.rh = cur.y1 - cur.y0;
		a.sw = low.x1 - low.x0;
		a.sh = low.y1 - low.y0;
		a.px = cur.x0 & 1;
		a.py = cur.y0 & 1;
		if(a.rw == 0 || a.rh == 0)
			break; // nothing below an empty resolution
		const bool lowest = (res == 1);
		for(int c = 0; c < g.ncomp; ++c)
		{
			a.pl[c].full = (void*)src[c];
			a.pl[c].f_stride = src_stride;
			a.pl[c].out_ll = lowest ? g.coeff[c] : g.scratch[c][ping];
			a.pl[c].o_stride = lowest ? g.coeff_stride : g.scratch_stride;
			a.pl[c].out_bands = g.coeff[c];
			a.pl[c].ob_stride = g.coeff_stride;
		}
		int lfin = first ? fin : 0;
		for(int c = 0; c < 3; ++c)
			a.shift[c] = shift ? shift[c < g.ncomp ? c : 0] : 0;
		if(g.ncomp == 3)
		{
			if(lfin == 1)
			{
				grkb_set_error("fdwt: per-component DC shift on a 3-component group is done by the caller");
				return -1;
			}
			if(grkb_launch_fdwt_level(a, reversible, 3, lfin, s))
				return -1;
		}
		else
		{
			for(int c = 0; c < g.ncomp; ++c)
			{
				DwtLevelArgs b = a;
				b.pl[0] = a.pl[c];
				b.shift[0] = a.shift[c];
				if(grkb_launch_fdwt_level(b, reversible, 1, lfin, s))
					return -1;
			}
		}
		for(int c = 0; c < g.ncomp; ++c)
			src[c] = a.pl[c].out_ll;
		src_stride = lowest ? g.coeff_stride : g.scratch_stride;
		ping ^= 1;
		first = false;
	}
	return 0;
}

// ------------------------------------------------------------------------------------------
// C-ABI, section A

namespace
{
struct DevBuf
{
	void* p = nullptr;
	~DevBuf()
	{
		if(p)
			cudaFree(p);
	}
	int alloc(size_t bytes)
	{
		if(bytes == 0)
			bytes = 16;
		cudaError_t e = cudaMalloc(&p, bytes);
		if(e != cudaSuccess)
		{
			grkb_set_error("cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
			p = nullptr;
			return -1;
		}
		return 0;
	}
};

// common driver for the four host wavelet entry points
int dwt_host_impl(void* const* planes, uint32_t ncomp, uint32_t stride, grkb_rect tc,
				  uint32_t numres, int reversible, int forward, int fin, const int32_t* shift,
				  const int32_t* lo, const int32_t* hi)
{
	if(ncomp != 1 && ncomp != 3)
	{
		grkb_set_error("ncomp must be 1 or 3 for a geometry group");
		return -1;
	}
	if(tc.x1 < tc.x0 || tc.y1 < tc.y0 || numres == 0 || numres > 33)
	{
		grkb_set_error("bad geometry");
		return -1;
	}
	const uint32_t W = tc.x1 - tc.x0, Hh = tc.y1 - tc.y0;
	if(W == 0 || Hh == 0)
		return 0;
	if(stride < W)
	{
		grkb_set_error("stride smaller than width");
		return -1;
	}
	const uint32_t dstride = grkb_stride(W);
	const size_t plane_bytes = (size_t)dstride * Hh * 4;
	uint32_t sstride = 32;
	const size_t sbytes = grkb_dwt_scratch_bytes(tc, numres, &sstride);
	DevBuf coeff[3], image[3], scr[3][2];
	DwtGroup g{};
	g.ncomp = (int)ncomp;
	g.coeff_stride = dstride;
	g.image_stride = dstride;
	g.scratch_stride = sstride;
	for(uint32_t c = 0; c < ncomp; ++c)
	{
		if(coeff[c].alloc(plane_bytes) || image[c].alloc(plane_bytes) || scr[c][0].alloc(sbytes) ||
		   scr[c][1].alloc(sbytes))
			return -1;
		g.coeff[c] = coeff[c].p;
		g.image[c] = image[c].p;
		g.scratch[c][0] = scr[c][0].p;
		g.scratch[c][1] = scr[c][1].p;
		void* dst = forward ? image[c].p : coeff[c].p;
		GRKB_CUDA_CHECK(cudaMemcpy2D(dst, (size_t)dstride * 4, planes[c], (size_t)stride * 4,
									 (size_t)W * 4, Hh, cudaMemcpyHostToDevice));
	}
	int rc = forward ? grkb_fdwt_device(g, tc, numres, reversible, fin, shift, 0)
					 : grkb_idwt_device(g, tc, numres, reversible, fin, shift, lo, hi, 0);
	if(rc)
		return rc;
	GRKB_CUDA_CHECK(cudaStreamSynchronize(0));
	for(uint32_t c = 0; c < ncomp; ++c)
	{
		void* src = forward ? coeff[c].p : image[c].p;
		GRKB_CUDA_CHECK(cudaMemcpy2D(planes[c], (size_t)stride * 4, src, (size_t)dstride * 4,
									 (size_t)W * 4, Hh, cudaMemcpyDeviceToHost));
	}
	return 0;
}
} // namespace

extern "C" {

int grkb_version(void)
{
	return GRKB_VERSION;
}

int grkb_device_count(void)
{
	int n = 0;
	if(cudaGetDeviceCount(&n) != cudaSuccess)
	{
		cudaGetLastError();
		return 0;
	}
	return n;
}

int grkb_set_device(int device)
{
	GRKB_CUDA_CHECK(cudaSetDevice(device));
	return 0;
}

const char* grkb_last_error(void)
{
	return g_err;
}

// Restrict the calling thread to the CPUs next to `device` (its PCI function's local_cpulist), so
// that the page-locked buffers it allocates afterwards are first touched -- and the copies to and
// from them stay -- on the GPU's own NUMA node.  With one process (or thread) per GPU on a
// two-socket box the D2H of the int32 planes otherwise crosses the socket link for half the GPUs.
// Returns 0 when the affinity was set, 1 when the box gives no topology (nothing changed).
int grkb_bind_host_to_device(int device)
{
	char bus[32] = {};
	GRKB_CUDA_CHECK(cudaDeviceGetPCIBusId(bus, (int)sizeof(bus), device));
	for(char* c = bus; *c; ++c)
		*c = (char)tolower(*c);
	char path[128];
	snprintf(path, sizeof(path), "/sys/bus/pci/devices/%s/local_cpulist", bus);
	FILE* f = fopen(path, "r");
	if(!f)
		return 1;
	char line[4096] = {};
	const bool got = fgets(line, sizeof(line), f) != nullptr;
	fclose(f);
	if(!got)
		return 1;
	cpu_set_t allowed, want;
	CPU_ZERO(&allowed);
	CPU_ZERO(&want);
	if(sched_getaffinity(0, sizeof(allowed), &allowed) != 0)
		return 1;
	int n = 0;
	for(char* tok = strtok(line, ",\n"); tok; tok = strtok(nullptr, ",\n"))
	{
		int a = 0, b = 0;
		const int k = sscanf(tok, "%d-%d", &a, &b);
		if(k < 1)
			continue;
		if(k == 1)
			b = a;
		for(int c = a; c <= b && c < CPU_SETSIZE; ++c)
			if(c >= 0 && CPU_ISSET(c, &allowed))
			{
				CPU_SET(c, &want);
				++n;
			}
	}
	if(n == 0 || n == CPU_COUNT(&allowed))
		return 1; // no local CPUs among the allowed ones, or nothing to narrow
	return sched_setaffinity(0, sizeof(want), &want) == 0 ? 0 : 1;
}

int grkb_dwt_host(void* data, uint32_t stride, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1,
				  uint32_t numres, int reversible, int forward)
{
	void* planes[1] = {data};
	int32_t z[3] = {0, 0, 0};
	return dwt_host_impl(planes, 1, stride, grkb_rect{x0, y0, x1, y1}, numres, reversible, forward,
						 0, z, z, z);
}

int grkb_idwt_mct_host(void* const* planes, uint32_t ncomp, uint32_t stride, uint32_t x0,
					   uint32_t y0, uint32_t x1, uint32_t y1, uint32_t numres, int reversible,
					   int mct, const int32_t* shift, const int32_t* lo, const int32_t* hi)
{
	const grkb_rect tc{x0, y0, x1, y1};
	if(mct)
	{
		if(ncomp != 3)
		{
			grkb_set_error("MCT needs exactly 3 components");
			return -1;
		}
		return dwt_host_impl(planes, 3, stride, tc, numres, reversible, 0, 2, shift, lo, hi);
	}
	for(uint32_t c = 0; c < ncomp; ++c)
	{
		void* one[1] = {planes[c]};
		int32_t s[3] = {shift[c], 0, 0}, l[3] = {lo[c], 0, 0}, h[3] = {hi[c], 0, 0};
		if(dwt_host_impl(one, 1, stride, tc, numres, reversible, 0, 1, s, l, h))
			return -1;
	}
	return 0;
}

int grkb_mct_fdwt_host(void* const* planes, uint32_t ncomp, uint32_t stride, uint32_t x0,
					   uint32_t y0, uint32_t x1, uint32_t y1, uint32_t numres, int reversible,
					   int mct, const int32_t* shift)
{
	const grkb_rect tc{x0, y0, x1, y1};
	int32_t z[3] = {0, 0, 0};
	if(mct)
	{
		if(ncomp != 3)
		{
			grkb_set_error("MCT needs exactly 3 components");
			return -1;
		}
		return dwt_host_impl(planes, 3, stride, tc, numres, reversible, 1, 2, shift, z, z);
	}
	for(uint32_t c = 0; c < ncomp; ++c)
	{
		void* one[1] = {planes[c]};
		int32_t s[3] = {shift[c], 0, 0};
		if(dwt_host_impl(one, 1, stride, tc, numres, reversible, 1, 1, s, z, z))
			return -1;
	}
	return 0;
}

int grkb_mct_host(void* c0, void* c1, void* c2, uint64_t n, int mode, const int32_t* shift,
				  const int32_t* lo, const int32_t* hi)
{
	if(n == 0)
		return 0;
	void* h[3] = {c0, c1, c2};
	DevBuf d[3];
	MctArgs m{};
	m.n = n;
	for(int c = 0; c < 3; ++c)
	{
		if(d[c].alloc(n * 4))
			return -1;
		GRKB_CUDA_CHECK(cudaMemcpy(d[c].p, h[c], n * 4, cudaMemcpyHostToDevice));
		m.c[c] = d[c].p;
		m.shift[c] = shift ? shift[c] : 0;
		m.lo[c] = lo ? lo[c] : 0;
		m.hi[c] = hi ? hi[c] : 0;
	}
	if(grkb_launch_mct(m, mode, 0))
		return -1;
	for(int c = 0; c < 3; ++c)
		GRKB_CUDA_CHECK(cudaMemcpy(h[c], d[c].p, n * 4, cudaMemcpyDeviceToHost));
	return 0;
}

int grkb_dc_shift_host(void* c, uint64_t n, int mode, int32_t shift, int32_t lo, int32_t hi)
{
	if(n == 0)
		return 0;
	DevBuf d;
	if(d.alloc(n * 4))
		return -1;
	GRKB_CUDA_CHECK(cudaMemcpy(d.p, c, n * 4, cudaMemcpyHostToDevice));
	MctArgs m{};
	m.n = n;
	m.c[0] = d.p;
	m.shift[0] = shift;
	m.lo[0] = lo;
	m.hi[0] = hi;
	if(grkb_launch_dc(m, mode, 0))
		return -1;
	GRKB_CUDA_CHECK(cudaMemcpy(c, d.p, n * 4, cudaMemcpyDeviceToHost));
	return 0;
}

} // extern "C"

// ------------------------------------------------------------------------------------------
// C-ABI, section C: device-resident timing of the wavelet back/front end (bench.py roofline)

extern "C" int grkb_bench_dwt(uint32_t w, uint32_t h, uint32_t numres, int reversible, int forward,
							  uint32_t ncomp, int mct, uint32_t iters, uint32_t warmup,
							  float* ms_per_iter)
{
	if((ncomp != 1 && ncomp != 3) || w == 0 || h == 0 || iters == 0)
	{
		grkb_set_error("grkb_bench_dwt: bad arguments");
		return -1;
	}
	const grkb_rect tc{0, 0, w, h};
	const uint32_t dstride = grkb_stride(w);
	const size_t plane_bytes = (size_t)dstride * h * 4;
	uint32_t sstride = 32;
	const size_t sbytes = grkb_dwt_scratch_bytes(tc, numres, &sstride);
	DevBuf coeff[3], image[3], scr[3][2];
	DwtGroup g{};
	g.ncomp = (int)ncomp;
	g.coeff_stride = g.image_stride = dstride;
	g.scratch_stride = sstride;
	std::vector<int32_t> host((size_t)dstride * h);
	for(uint32_t c = 0; c < ncomp; ++c)
	{
		if(coeff[c].alloc(plane_bytes) || image[c].alloc(plane_bytes) || scr[c][0].alloc(sbytes) ||
		   scr[c][1].alloc(sbytes))
			return -1;
		uint32_t seed = 12345u + c;
		for(size_t i = 0; i < host.size(); ++i)
		{
			seed = seed * 1664525u + 1013904223u;
			int32_t v = (int32_t)(seed >> 20) - 2048;
			if(!reversible && !forward)
			{
				float f = (float)v;
				memcpy(&host[i], &f, 4);
			}
			else
				host[i] = forward ? (v + 2048) : v;
		}
		GRKB_CUDA_CHECK(cudaMemcpy(forward ? image[c].p : coeff[c].p, host.data(), plane_bytes,
								   cudaMemcpyHostToDevice));
		g.coeff[c] = coeff[c].p;
		g.image[c] = image[c].p;
		g.scratch[c][0] = scr[c][0].p;
		g.scratch[c][1] = scr[c][1].p;
	}
	int32_t shift[3] = {2048, 2048, 2048}, lo[3] = {0, 0, 0}, hi[3] = {4095, 4095, 4095};
	const int fin = (mct && ncomp == 3) ? 2 : 0;
	cudaEvent_t e0, e1;
	GRKB_CUDA_CHECK(cudaEventCreate(&e0));
	GRKB_CUDA_CHECK(cudaEventCreate(&e1));
	for(uint32_t it = 0; it < warmup + iters; ++it)
	{
		if(it == warmup)
			GRKB_CUDA_CHECK(cudaEventRecord(e0, 0));
		int rc = forward ? grkb_fdwt_device(g, tc, numres, reversible, fin, shift, 0)
						 : grkb_idwt_device(g, tc, numres, reversible, fin, shift, lo, hi, 0);
		if(rc)
			return rc;
	}
	GRKB_CUDA_CHECK(cudaEventRecord(e1, 0));
	GRKB_CUDA_CHECK(cudaEventSynchronize(e1));
	float ms = 0;
	GRKB_CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
	*ms_per_iter = ms / (float)iters;
	cudaEventDestroy(e0);
	cudaEventDestroy(e1);
	return 0;
}


// Host <-> device link bandwidth with page-locked memory, measured the way the decoders use the
// link: one cudaMemcpyAsync per direction on a stream of its own.  h2d / d2h: each direction alone;
// duplex[0..1]: both at once (what a pipelined decode sees).  GB/s (1e9 bytes).
extern "C" int grkb_bench_link(size_t bytes, uint32_t iters, double* h2d_gbs, double* d2h_gbs, double duplex_gbs[2])
{
	if(bytes < 4096 || iters == 0)
	{
		grkb_set_error("grkb_bench_link: bad arguments");
		return -1;
	}
	void *hin = nullptr, *hout = nullptr;
	DevBuf din, dout;
	if(cudaMallocHost(&hin, bytes) != cudaSuccess || cudaMallocHost(&hout, bytes) != cudaSuccess || din.alloc(bytes) ||
	   dout.alloc(bytes))
	{
		if(hin)
			cudaFreeHost(hin);
		if(hout)
			cudaFreeHost(hout);
		grkb_set_error("grkb_bench_link: allocation failed");
		return -1;
	}
	memset(hin, 1, bytes);
	memset(hout, 0, bytes);
	cudaStream_t s0, s1;
	cudaEvent_t e[4];
	int rc = -1;
	if(cudaStreamCreateWithFlags(&s0, cudaStreamNonBlocking) == cudaSuccess &&
	   cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking) == cudaSuccess)
	{
		for(auto& x : e)
			cudaEventCreate(&x);
		auto run = [&](bool up, bool down, float& ms_up, float& ms_down) {
			// warm-up
			if(up)
				cudaMemcpyAsync(din.p, hin, bytes, cudaMemcpyHostToDevice, s0);
			if(down)
				cudaMemcpyAsync(hout, dout.p, bytes, cudaMemcpyDeviceToHost, s1);
			cudaStreamSynchronize(s0);
			cudaStreamSynchronize(s1);
			cudaEventRecord(e[0], s0);
			cudaEventRecord(e[2], s1);
			for(uint32_t i = 0; i < iters; ++i)
			{
				if(up)
					cudaMemcpyAsync(din.p, hin, bytes, cudaMemcpyHostToDevice, s0);
				if(down)
					cudaMemcpyAsync(hout, dout.p, bytes, cudaMemcpyDeviceToHost, s1);
			}
			cudaEventRecord(e[1], s0);
			cudaEventRecord(e[3], s1);
			cudaStreamSynchronize(s0);
			cudaStreamSynchronize(s1);
			cudaEventElapsedTime(&ms_up, e[0], e[1]);
			cudaEventElapsedTime(&ms_down, e[2], e[3]);
		};
		float a = 0, b = 0, c = 0, d = 0, x = 0;
		run(true, false, a, x);
		run(false, true, x, b);
		run(true, true, c, d);
		const double gb = (double)bytes * iters / 1e9;
		if(h2d_gbs)
			*h2d_gbs = a > 0 ? gb / (a * 1e-3) : 0;
		if(d2h_gbs)
			*d2h_gbs = b > 0 ? gb / (b * 1e-3) : 0;
		if(duplex_gbs)
		{
			duplex_gbs[0] = c > 0 ? gb / (c * 1e-3) : 0;
			duplex_gbs[1] = d > 0 ? gb / (d * 1e-3) : 0;
		}
		rc = cudaGetLastError() == cudaSuccess ? 0 : -1;
		for(auto& x2 : e)
			cudaEventDestroy(x2);
		cudaStreamDestroy(s0);
		cudaStreamDestroy(s1);
	}
	cudaFreeHost(hin);
	cudaFreeHost(hout);
	return rc;
}
