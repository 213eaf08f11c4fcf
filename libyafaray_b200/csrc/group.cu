// yafgpu_group: one scene on N GPUs of one box behind the C ABI (include/yafgpu.h "multi-GPU").
//
// The reference renders a frame with a pool of worker threads pulling tiles from ImageSplitter (imagesplitter.cc:37-50) and summing
// into one ImageFilm; distributed rendering is done by summing the accumulator films of several nodes (imagefilm.cc:1506-1535).  The
// group does both on one 8-GPU box: the scene is replicated (2.1 GB at 10^7 triangles), member k of N renders the raster tiles t with
// t mod N = k of every pass (yafgpu_pass_desc.part_index / part_count, one host thread per GPU), each into its own full-size film —
// filter footprints cross tile borders — and after the pass the N films are summed on the first device: peer copies over NVLink
// (cudaMemcpyPeerAsync, staged by the driver where peer access is not available) and one add kernel each, queued on that device's
// stream behind events of the members' streams, so the host does not wait for it and the members' next pass only waits for the copy
// that reads their film.  No collective runs on the data path: all randomness is a function of (pixel, sample, tree position), so
// the sum equals the single-GPU film up to float summation order.
#include "context.h"
#include <thread>
#include <vector>
#include <string>

struct yafgpu_group
{
	std::vector<yafgpu_ctx *> ctx;
	std::vector<int> devices;
	std::string err;
	float *red = nullptr;      // reduced film on devices[0] (N > 1)
	float *tmp = nullptr;      // staging for one peer film on devices[0]
	size_t film_floats = 0;
	std::vector<cudaEvent_t> pass_done;   // per member: recorded on its stream after its part of the pass
	cudaEvent_t reduce_done = nullptr;    // recorded on the first member's stream after the sum; members wait for it before their next pass
	bool reduced_valid = false;
	uint64_t reduces = 0; double ms_reduce = 0.0;
	cudaEvent_t ev_r0 = nullptr, ev_r1 = nullptr;
};

static std::string g_group_create_err;

namespace yg {

__global__ void __launch_bounds__(256) k_film_sum(float4 *__restrict__ dst, const float4 *__restrict__ src, size_t n4, float *__restrict__ dst_tail, const float *__restrict__ src_tail, int n_tail)
{
	for(size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x)
	{
		float4 a = dst[i]; const float4 b = src[i];
		a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
		dst[i] = a;
	}
	if(blockIdx.x == 0 && (int)threadIdx.x < n_tail) dst_tail[threadIdx.x] += src_tail[threadIdx.x];
}

static int gErr(yafgpu_group *g, int code, const std::string &msg) { if(g) g->err = msg; else g_group_create_err = msg; return code; }

// run f(k) for every member on its own host thread; first failure wins
template<class F> static int forAll(yafgpu_group *g, const char *what, F f)
{
	const size_t n = g->ctx.size();
	std::vector<int> rc(n, YAFGPU_OK);
	if(n == 1) rc[0] = f(0);
	else
	{
		std::vector<std::thread> th;
		for(size_t k = 0; k < n; ++k) th.emplace_back([&, k]() { rc[k] = f((int)k); });
		for(std::thread &t : th) t.join();
	}
	for(size_t k = 0; k < n; ++k)
		if(rc[k] != YAFGPU_OK) return gErr(g, rc[k], std::string(what) + " on device " + std::to_string(g->devices[k]) + ": " + yafgpu_last_error(g->ctx[k]));
	return YAFGPU_OK;
}

static void freeReduce(yafgpu_group *g)
{
	if(g->ctx.empty()) return;
	cudaSetDevice(g->devices[0]);
	if(g->red) cudaFree(g->red);
	if(g->tmp) cudaFree(g->tmp);
	g->red = g->tmp = nullptr; g->film_floats = 0; g->reduced_valid = false;
}

// red = sum of the members' films, on the first member's stream (asynchronous to the host)
static int reduceFilms(yafgpu_group *g)
{
	const size_t n = g->ctx.size();
	if(n == 1) return YAFGPU_OK;
	yafgpu_ctx *c0 = g->ctx[0];
	cudaError_t e = cudaSetDevice(g->devices[0]);
	const size_t bytes = g->film_floats * sizeof(float);
	if(e == cudaSuccess) e = cudaEventRecord(g->ev_r0, c0->stream);
	if(e == cudaSuccess) e = cudaMemcpyAsync(g->red, c0->film.pix, bytes, cudaMemcpyDeviceToDevice, c0->stream);
	const size_t n4 = g->film_floats / 4; const int tail = (int)(g->film_floats - 4 * n4);
	for(size_t k = 1; k < n && e == cudaSuccess; ++k)
	{
		e = cudaStreamWaitEvent(c0->stream, g->pass_done[k], 0);
		if(e == cudaSuccess) e = cudaMemcpyPeerAsync(g->tmp, g->devices[0], g->ctx[k]->film.pix, g->devices[k], bytes, c0->stream);
		if(e == cudaSuccess)
		{
			k_film_sum<<<148 * 4, 256, 0, c0->stream>>>((float4 *)g->red, (const float4 *)g->tmp, n4, g->red + 4 * n4, g->tmp + 4 * n4, tail);
			e = cudaGetLastError();
		}
	}
	if(e == cudaSuccess) e = cudaEventRecord(g->ev_r1, c0->stream);
	if(e == cudaSuccess) e = cudaEventRecord(g->reduce_done, c0->stream);
	if(e != cudaSuccess) return gErr(g, YAFGPU_ERR_CUDA, std::string("film reduction: ") + cudaGetErrorString(e));
	g->reduced_valid = true; g->reduces++;
	return YAFGPU_OK;
}

} // namespace yg

using namespace yg;

extern "C" {

const char *yafgpu_group_last_error(const yafgpu_group *g) { return g ? g->err.c_str() : g_group_create_err.c_str(); }
int yafgpu_group_size(const yafgpu_group *g) { return g ? (int)g->ctx.size() : 0; }
yafgpu_ctx *yafgpu_group_ctx(yafgpu_group *g, int i) { return (g && i >= 0 && (size_t)i < g->ctx.size()) ? g->ctx[i] : nullptr; }

void yafgpu_group_destroy(yafgpu_group *g)
{
	if(!g) return;
	for(size_t k = 0; k < g->ctx.size(); ++k) { cudaSetDevice(g->devices[k]); if(g->ctx[k]->stream) cudaStreamSynchronize(g->ctx[k]->stream); }
	freeReduce(g);
	if(!g->ctx.empty())
	{
		cudaSetDevice(g->devices[0]);
		if(g->reduce_done) cudaEventDestroy(g->reduce_done);
		if(g->ev_r0) cudaEventDestroy(g->ev_r0);
		if(g->ev_r1) cudaEventDestroy(g->ev_r1);
	}
	for(size_t k = 0; k < g->ctx.size(); ++k)
	{
		cudaSetDevice(g->devices[k]);
		if(k < g->pass_done.size() && g->pass_done[k]) cudaEventDestroy(g->pass_done[k]);
		yafgpu_destroy(g->ctx[k]);
	}
	delete g;
}

int yafgpu_group_create(yafgpu_group **out, const int *devices, int n_devices, unsigned flags)
{
	if(!out) return gErr(nullptr, YAFGPU_ERR_INVALID, "yafgpu_group_create: out is NULL");
	*out = nullptr;
	int n_dev = 0;
	cudaError_t e = cudaGetDeviceCount(&n_dev);
	if(e != cudaSuccess || n_dev == 0)
		return gErr(nullptr, YAFGPU_ERR_CUDA, std::string("yafgpu_group_create: no usable CUDA device (") + cudaGetErrorString(e) + "); there is no CPU fallback");
	std::vector<int> devs;
	if(!devices || n_devices <= 0) for(int k = 0; k < n_dev; ++k) devs.push_back(k);   // all GPUs of the box
	else devs.assign(devices, devices + n_devices);
	for(size_t a = 0; a < devs.size(); ++a)
	{
		if(devs[a] < 0 || devs[a] >= n_dev) return gErr(nullptr, YAFGPU_ERR_INVALID, "yafgpu_group_create: device index out of range");
		for(size_t b = 0; b < a; ++b) if(devs[a] == devs[b] && !(flags & YAFGPU_FLAG_SHARE_DEVICE)) return gErr(nullptr, YAFGPU_ERR_INVALID, "yafgpu_group_create: a device is listed twice");
	}
	yafgpu_group *g = new yafgpu_group();
	g->devices = devs;
	for(int d : devs)
	{
		yafgpu_ctx *c = nullptr;
		const int rc = yafgpu_create(&c, d, flags & ~(unsigned)YAFGPU_FLAG_SHARE_DEVICE);
		if(rc != YAFGPU_OK) { const std::string m = yafgpu_last_error(nullptr); yafgpu_group_destroy(g); return gErr(nullptr, rc, m); }
		g->ctx.push_back(c);
		cudaEvent_t ev = nullptr;
		cudaSetDevice(d);
		cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
		g->pass_done.push_back(ev);
	}
	cudaSetDevice(devs[0]);
	cudaEventCreateWithFlags(&g->reduce_done, cudaEventDisableTiming);
	cudaEventCreate(&g->ev_r0); cudaEventCreate(&g->ev_r1);
	// direct NVLink copies where the topology allows them (cudaMemcpyPeerAsync is staged through the host otherwise)
	for(size_t k = 1; k < devs.size(); ++k)
	{
		int can = 0;
		if(devs[k] == devs[0]) continue;
		if(cudaDeviceCanAccessPeer(&can, devs[0], devs[k]) == cudaSuccess && can) { cudaSetDevice(devs[0]); cudaDeviceEnablePeerAccess(devs[k], 0); }
		if(cudaDeviceCanAccessPeer(&can, devs[k], devs[0]) == cudaSuccess && can) { cudaSetDevice(devs[k]); cudaDeviceEnablePeerAccess(devs[0], 0); }
	}
	cudaGetLastError(); // "peer access already enabled" is not an error here
	*out = g;
	return YAFGPU_OK;
}

int yafgpu_group_upload_scene(yafgpu_group *g, const yafgpu_scene_desc *desc)
{
	if(!g || !desc) return gErr(g, YAFGPU_ERR_INVALID, "yafgpu_group_upload_scene: NULL argument");
	freeReduce(g);
	int rc = forAll(g, "scene upload", [&](int k) { return yafgpu_upload_scene(g->ctx[k], desc); });
	if(rc != YAFGPU_OK) return rc;
	g->film_floats = (size_t)5 * g->ctx[0]->film.w * g->ctx[0]->film.h;
	if(g->ctx.size() > 1)
	{
		cudaError_t e = cudaSetDevice(g->devices[0]);
		if(e == cudaSuccess) e = cudaMalloc(&g->red, g->film_floats * sizeof(float));
		if(e == cudaSuccess) e = cudaMalloc(&g->tmp, g->film_floats * sizeof(float));
		if(e != cudaSuccess) return gErr(g, YAFGPU_ERR_CUDA, std::string("reduced film allocation: ") + cudaGetErrorString(e));
	}
	return YAFGPU_OK;
}

int yafgpu_group_film_clear(yafgpu_group *g)
{
	if(!g) return YAFGPU_ERR_INVALID;
	g->reduced_valid = false;
	return forAll(g, "film clear", [&](int k) { return yafgpu_film_clear(g->ctx[k]); });
}

int yafgpu_group_write_film(yafgpu_group *g, const float *rgbaw)
{
	if(!g || !rgbaw) return gErr(g, YAFGPU_ERR_INVALID, "NULL argument");
	g->reduced_valid = false;
	// the films are summed: the first member takes the loaded accumulators, the others start from zero
	return forAll(g, "film upload", [&](int k) { return k == 0 ? yafgpu_write_film(g->ctx[0], rgbaw) : yafgpu_film_clear(g->ctx[k]); });
}

int yafgpu_group_render_pass(yafgpu_group *g, const yafgpu_pass_desc *pass)
{
	if(!g || !pass) return gErr(g, YAFGPU_ERR_INVALID, "NULL argument");
	const int n = (int)g->ctx.size();
	int rc = forAll(g, "render pass", [&](int k)
	{
		yafgpu_ctx *c = g->ctx[k];
		cudaSetDevice(g->devices[k]);
		// this member's film must not change before the previous reduction has copied it
		if(n > 1 && g->reduced_valid) cudaStreamWaitEvent(c->stream, g->reduce_done, 0);
		yafgpu_pass_desc p = *pass;
		if(n > 1) { p.part_index = k; p.part_count = n; }
		const int r = yafgpu_render_pass(c, &p);
		if(r == YAFGPU_OK && n > 1) cudaEventRecord(g->pass_done[k], c->stream);
		return r;
	});
	if(rc != YAFGPU_OK) return rc;
	return reduceFilms(g);
}

int yafgpu_group_read_film(yafgpu_group *g, float *out)
{
	if(!g || !out) return gErr(g, YAFGPU_ERR_INVALID, "NULL argument");
	if(g->ctx.size() == 1)
	{
		const int rc = yafgpu_read_film(g->ctx[0], out);
		return rc == YAFGPU_OK ? rc : gErr(g, rc, yafgpu_last_error(g->ctx[0]));
	}
	if(!g->red) return gErr(g, YAFGPU_ERR_STATE, "no scene uploaded");
	if(!g->reduced_valid) { const int rc = reduceFilms(g); if(rc != YAFGPU_OK) return rc; }
	yafgpu_ctx *c0 = g->ctx[0];
	cudaError_t e = cudaSetDevice(g->devices[0]);
	if(e == cudaSuccess) e = cudaMemcpyAsync(out, g->red, g->film_floats * sizeof(float), cudaMemcpyDeviceToHost, c0->stream);
	if(e == cudaSuccess) e = cudaStreamSynchronize(c0->stream);
	if(e != cudaSuccess) return gErr(g, YAFGPU_ERR_CUDA, std::string("reduced film read-back: ") + cudaGetErrorString(e));
	float ms = 0.f;
	if(cudaEventElapsedTime(&ms, g->ev_r0, g->ev_r1) == cudaSuccess) g->ms_reduce = ms;
	return YAFGPU_OK;
}

int yafgpu_group_next_pass(yafgpu_group *g, const yafgpu_aa_noise_params *prm, int64_t *n_resample)
{
	if(!g || !prm) return gErr(g, YAFGPU_ERR_INVALID, "NULL argument");
	if(g->ctx.size() == 1)
	{
		const int rc = yafgpu_next_pass(g->ctx[0], prm, n_resample);
		return rc == YAFGPU_OK ? rc : gErr(g, rc, yafgpu_last_error(g->ctx[0]));
	}
	if(!g->red) return gErr(g, YAFGPU_ERR_STATE, "no scene uploaded");
	if(!g->reduced_valid) { const int rc = reduceFilms(g); if(rc != YAFGPU_OK) return rc; }
	// the decision is taken once, on the summed film (first device, behind the film sum on its stream); every member gets the same mask
	yafgpu_ctx *c0 = g->ctx[0];
	cudaSetDevice(g->devices[0]);
	const int rc = nextPassMask(c0, g->red, prm, c0->aa_mask, n_resample);
	if(rc != YAFGPU_OK) return gErr(g, rc, yafgpu_last_error(c0));
	for(size_t k = 0; k < g->ctx.size(); ++k) { if(k) g->ctx[k]->aa_mask = c0->aa_mask; g->ctx[k]->aa_mask_valid = true; }
	return YAFGPU_OK;
}

int yafgpu_group_sync(yafgpu_group *g)
{
	if(!g) return YAFGPU_ERR_INVALID;
	return forAll(g, "sync", [&](int k) { return yafgpu_sync(g->ctx[k]); });
}

int yafgpu_group_set_aa_multipliers(yafgpu_group *g, float light_mult, float indirect_mult)
{
	if(!g) return YAFGPU_ERR_INVALID;
	return forAll(g, "set_aa_multipliers", [&](int k) { return yafgpu_set_aa_multipliers(g->ctx[k], light_mult, indirect_mult); });
}

int yafgpu_group_set_option(yafgpu_group *g, const char *name, long long value)
{
	if(!g) return YAFGPU_ERR_INVALID;
	return forAll(g, "set_option", [&](int k) { return yafgpu_set_option(g->ctx[k], name, value); });
}

int yafgpu_group_get_stats(yafgpu_group *g, yafgpu_stats *out)
{
	if(!g || !out) return gErr(g, YAFGPU_ERR_INVALID, "NULL argument");
	memset(out, 0, sizeof(*out));
	for(size_t k = 0; k < g->ctx.size(); ++k)
	{
		yafgpu_stats s;
		const int rc = yafgpu_get_stats(g->ctx[k], &s);
		if(rc != YAFGPU_OK) return gErr(g, rc, yafgpu_last_error(g->ctx[k]));
		if(k == 0) *out = s;
		else
		{	// counts add up over the members, times are the slowest member's
			out->rays_camera += s.rays_camera; out->rays_extension += s.rays_extension; out->rays_shadow += s.rays_shadow; out->rays_shadow_transp += s.rays_shadow_transp;
			out->rays_light_probe += s.rays_light_probe; out->camera_samples += s.camera_samples; out->kernel_launches += s.kernel_launches;
			out->nodes_visited += s.nodes_visited; out->tris_tested += s.tris_tested; out->nodes_visited_shadow += s.nodes_visited_shadow; out->tris_tested_shadow += s.tris_tested_shadow;
			out->waves += s.waves; out->wave_retries += s.wave_retries; out->nonfinite_samples += s.nonfinite_samples;
			out->ms_total = s.ms_total > out->ms_total ? s.ms_total : out->ms_total; out->ms_build = s.ms_build > out->ms_build ? s.ms_build : out->ms_build;
		}
	}
	out->film_reduces = g->reduces; out->ms_film_reduce = g->ms_reduce;
	return YAFGPU_OK;
}

} // extern "C"
