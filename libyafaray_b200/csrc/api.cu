// C ABI of include/yafgpu.h: context life cycle, scene upload (Scene::update() geometry branch, scene.cc:788-831),
// batch tracing (Scene::intersect / isShadowed probes) and film access.  The integrator lives in integrator.cu.
#include "context.h"
#include "host_prep.h"
#include "bvh_build.h"
#include "dev_pool.h"
#include "traverse.cuh"
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <new>
#include <mutex>
#include <thread>
#include <chrono>

#define YG_MAX_DEVICES 64 /* per-device tables of yafgpu_trace_static */

static std::string g_create_err;

namespace yg {

int setErr(yafgpu_ctx *ctx, int code, const std::string &msg)
{
	if(ctx) ctx->err = msg; else g_create_err = msg;
	return code;
}
int cudaErr(yafgpu_ctx *ctx, cudaError_t e, const char *what)
{
	return setErr(ctx, YAFGPU_ERR_CUDA, std::string("CUDA error: ") + cudaGetErrorString(e) + " in " + what);
}

int faultCheck(yafgpu_ctx *c, bool fetch)
{
	if(fetch)
	{
		cudaError_t e = cudaMemcpyAsync(c->h_fault, c->d_fault, sizeof(int), cudaMemcpyDeviceToHost, c->stream);
		if(e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
		if(e != cudaSuccess) return cudaErr(c, e, "traversal fault word");
	}
	if(*c->h_fault == 0) return YAFGPU_OK;
	*c->h_fault = 0;
	cudaMemsetAsync(c->d_fault, 0, sizeof(int), c->stream);
	return setErr(c, YAFGPU_ERR_CUDA, "BVH traversal stack overflow: a subtree was skipped, the result is discarded (tree deeper than the traversal stack)");
}

template<class T> static cudaError_t upload(T *&dst, const T *src, size_t n, cudaStream_t s)
{
	dst = nullptr;
	cudaError_t e = poolAlloc(&dst, sizeof(T) * n, s);
	if(e != cudaSuccess) return e;
	if(n) e = cudaMemcpyAsync(dst, src, sizeof(T) * n, cudaMemcpyHostToDevice, s);
	return e;
}

static void freeScene(yafgpu_ctx *c)
{
	poolFree(c->d_points, c->stream); poolFree(c->d_normals, c->stream); poolFree(c->d_uv, c->stream); poolFree(c->d_tri_ng, c->stream); poolFree(c->d_tri_vidx, c->stream);
	poolFree(c->d_tri_mat, c->stream); poolFree(c->d_mat_triflags, c->stream); poolFree(c->d_tris, c->stream); poolFree(c->d_objects, c->stream); poolFree(c->d_mats, c->stream);
	poolFree(c->d_lights, c->stream); poolFree(c->d_light_tables, c->stream); poolFree(c->d_bg_tex, c->stream); poolFree(c->d_nodes, c->stream); poolFree(c->d_acc_tris, c->stream);
	poolFree(c->film.pix, c->stream); poolFree(c->d_mat_bucket, c->stream); c->d_mat_bucket = nullptr; poolFree(c->d_light_accels, c->stream); c->d_light_accels = nullptr;
	for(void *q : c->light_bvh_ptrs) poolFree(q, c->stream);
	c->light_bvh_ptrs.clear();
	for(void *q : c->tex_ptrs) poolFree(q, c->stream);
	c->tex_ptrs.clear();
	poolFree(c->d_cie, c->stream); c->d_cie = nullptr;
	poolFree(c->d_noise, c->stream); c->d_noise = nullptr;
	poolFree(c->d_textures, c->stream); c->d_textures = nullptr; poolFree(c->d_shader_nodes, c->stream); c->d_shader_nodes = nullptr;
	c->d_points = c->d_normals = c->d_uv = c->d_tri_ng = nullptr; c->d_tri_vidx = c->d_tri_mat = nullptr; c->d_mat_triflags = nullptr;
	c->d_tris = nullptr; c->d_objects = nullptr; c->d_mats = nullptr; c->d_lights = nullptr; c->d_light_tables = nullptr; c->d_bg_tex = nullptr;
	c->d_nodes = nullptr; c->d_acc_tris = nullptr; c->film.pix = nullptr;
	c->uploaded = false;
}

// builds TriShade records (global triangle order) from the flat index arrays
__global__ void k_tri_shade(int n, const int *__restrict__ vidx, const int *__restrict__ nidx, const int *__restrict__ uvidx,
                            const int *__restrict__ mat, const int *__restrict__ obj, const float *__restrict__ ng, TriShade *__restrict__ out)
{
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= n) return;
	TriShade t;
	for(int k = 0; k < 3; ++k)
	{
		t.v[k] = vidx[3 * i + k];
		t.n[k] = nidx ? nidx[3 * i + k] : -1;
		t.uv[k] = uvidx ? uvidx[3 * i + k] : -1;
		t.ng[k] = ng[3 * i + k];
	}
	t.mat = mat[i]; t.obj = obj[i];
	out[i] = t;
}

// ---- batch trace kernels (parity probe + microbench) -----------------------------------------------------------------
struct BatchIO
{
	const float4 *rays; float4 *hits; bool shadow;
	__device__ __forceinline__ void load(int i, V3 &o, V3 &d, float &tmin, float &dist) const
	{
		const float4 a = __ldg(rays + 2 * (size_t)i), b = __ldg(rays + 2 * (size_t)i + 1);
		o = mk(a.x, a.y, a.z); tmin = a.w; d = mk(b.x, b.y, b.z);
		dist = (b.w < 0.f) ? __int_as_float(0x7f800000) : b.w;
		if(shadow)
		{	// Scene::isShadowed scene.cc:962-970: shift the origin, shorten the segment
			o = o + d * a.w;
			if(!(b.w < 0.f)) dist = b.w - 2 * a.w;
		}
	}
	__device__ __forceinline__ void store(int i, const HitD &h, unsigned) const
	{
		hits[i] = shadow ? make_float4(h.tri >= 0 ? 1.f : 0.f, 0.f, 0.f, __int_as_float(h.tri)) : make_float4(h.t, h.u, h.v, __int_as_float(h.tri));
	}
};

// the warp-persistent traversal with ray regeneration the wavefront kernels use, on a caller-supplied ray array (at most 2^31 - 1 rays
// per call); `next` = the global fetch counter, zero at launch
template<bool COUNT>
__global__ void __launch_bounds__(128) k_trace_batch(Accel acc, const yafgpu_ray *__restrict__ rays, size_t n, yafgpu_hit *__restrict__ hits,
                                                        int mode, int *__restrict__ next, int refill_below, unsigned long long *__restrict__ counters)
{
	TravCounters tc; tc.nodes = 0; tc.tris = 0;
	__shared__ uint2 s_stack[YG_STACK_SHARED * 128];
	TravStack<YG_STACK_SHARED> st; bindSharedStack(st, s_stack);
	BatchIO io; io.rays = (const float4 *)rays; io.hits = (float4 *)hits; io.shadow = mode != YAFGPU_TRACE_CLOSEST;
	if(mode == YAFGPU_TRACE_CLOSEST) traversePersistent<false, COUNT>(acc, (int)n, next, io, st, &tc, refill_below);
	else traversePersistent<true, COUNT>(acc, (int)n, next, io, st, &tc, refill_below);
	if(COUNT) { atomicAdd(&counters[mode == YAFGPU_TRACE_CLOSEST ? CNT_NODES : CNT_NODES_SHADOW], tc.nodes); atomicAdd(&counters[mode == YAFGPU_TRACE_CLOSEST ? CNT_TRIS : CNT_TRIS_SHADOW], tc.tris); }
}

// ---- allocation-free batch trace (yafgpu_trace_static) ---------------------------------------------------------------------
// The batch trace needs device scratch: a copy of the caller's rays, room for the hit records and the ray-regeneration counter of
// traversePersistent.  yafgpu_trace takes it from the stream-ordered pool; callers that may not allocate get the same kernel over a
// file-scope array instead.  One array per GPU that loads the library, held until the process exits:
//   [0, 256)                                  int: next unclaimed ray (must be zero when the kernel starts)
//   [256, 256 + 32 n_max)                     yafgpu_ray[n_max]
//   [256 + 32 n_max, 256 + 48 n_max)          yafgpu_hit[n_max]
#define YG_STATIC_SCRATCH_BYTES (256 + (size_t)YAFGPU_TRACE_STATIC_MAX_RAYS * (sizeof(yafgpu_ray) + sizeof(yafgpu_hit)))
static __device__ __align__(256) unsigned char g_trace_static_scratch[YG_STATIC_SCRATCH_BYTES];

// Variant of k_trace_batch whose ray, hit and counter storage is carved out of one scratch block.  `scratch` is the DEVICE address of
// g_trace_static_scratch as cudaGetSymbolAddress reports it for the current device (staticScratch below).  In host code the symbol
// itself is only a stand-in, never an address a kernel may be given.
__global__ void __launch_bounds__(128) k_trace_batch_static(Accel acc, unsigned char *__restrict__ scratch, int n, int mode, int refill_below)
{
	TravCounters tc; tc.nodes = 0; tc.tris = 0;
	__shared__ uint2 s_stack[YG_STACK_SHARED * 128];
	TravStack<YG_STACK_SHARED> st; bindSharedStack(st, s_stack);
	int *next = (int *)scratch;
	BatchIO io;
	io.rays = (const float4 *)(scratch + 256);
	io.hits = (float4 *)(scratch + 256 + (size_t)YAFGPU_TRACE_STATIC_MAX_RAYS * sizeof(yafgpu_ray));
	io.shadow = mode != YAFGPU_TRACE_CLOSEST;
	if(mode == YAFGPU_TRACE_CLOSEST) traversePersistent<false, false>(acc, n, next, io, st, &tc, refill_below);
	else traversePersistent<true, false>(acc, n, next, io, st, &tc, refill_below);
}

} // namespace yg

using namespace yg;

extern "C" {

int yafgpu_abi_version(void) { return YAFGPU_ABI_VERSION; }

const char *yafgpu_last_error(const yafgpu_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

int yafgpu_create(yafgpu_ctx **out, int device, unsigned flags)
{
	if(!out) return setErr(nullptr, YAFGPU_ERR_INVALID, "yafgpu_create: out is NULL");
	*out = nullptr;
	int n_dev = 0;
	cudaError_t e = cudaGetDeviceCount(&n_dev);
	if(e != cudaSuccess || n_dev == 0)
		return setErr(nullptr, YAFGPU_ERR_CUDA, std::string("yafgpu_create: no usable CUDA device (") + cudaGetErrorString(e) + "); there is no CPU fallback");
	if(device < 0 || device >= n_dev) return setErr(nullptr, YAFGPU_ERR_INVALID, "yafgpu_create: device index out of range");
	if((e = cudaSetDevice(device)) != cudaSuccess) return cudaErr(nullptr, e, "cudaSetDevice");
	yafgpu_ctx *c = new(std::nothrow) yafgpu_ctx();
	if(!c) return setErr(nullptr, YAFGPU_ERR_NOMEM, "out of host memory");
	c->device = device; c->flags = flags;
	memset(&c->scene, 0, sizeof(c->scene)); memset(&c->film, 0, sizeof(c->film)); memset(&c->stats, 0, sizeof(c->stats));
	if((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess) { delete c; return cudaErr(nullptr, e, "cudaStreamCreate"); }
	cudaEventCreate(&c->ev0); cudaEventCreate(&c->ev1);
	for(int k = 0; k < 2; ++k) { cudaStreamCreateWithFlags(&c->side[k], cudaStreamNonBlocking); cudaEventCreateWithFlags(&c->ev_join[k], cudaEventDisableTiming); }
	cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming);
	{	// keep freed scene/builder blocks in the stream-ordered pool (dev_pool.h)
		cudaMemPool_t pool = nullptr;
		if(cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) { unsigned long long keep = ~0ull; cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep); }
	}
	FaureTables *ft = new FaureTables();
	yg_build_faure_tables(*ft);
	e = cudaMalloc(&c->d_faure, sizeof(FaureTables));
	if(e == cudaSuccess) e = cudaMemcpy(c->d_faure, ft, sizeof(FaureTables), cudaMemcpyHostToDevice);
	delete ft;
	if(e == cudaSuccess) e = cudaMalloc(&c->d_counters, sizeof(unsigned long long) * CNT_N);
	if(e == cudaSuccess) e = cudaMalloc(&c->d_fetch, sizeof(int));
	if(e == cudaSuccess) e = cudaMalloc(&c->d_fault, sizeof(int));
	if(e == cudaSuccess) e = cudaMemset(c->d_fault, 0, sizeof(int));
	if(e == cudaSuccess) e = cudaMallocHost(&c->h_fault, 2 * sizeof(int)); // [0] fault word mirror, [1] control word of yafgpu_trace_static
	if(e == cudaSuccess) { c->h_fault[0] = 0; c->h_fault[1] = 0; }
	if(e == cudaSuccess) e = cudaMemset(c->d_counters, 0, sizeof(unsigned long long) * CNT_N);
	if(e != cudaSuccess) { yafgpu_destroy(c); return cudaErr(nullptr, e, "context allocation"); }
	*out = c;
	return YAFGPU_OK;
}

void yafgpu_destroy(yafgpu_ctx *c)
{
	if(!c) return;
	cudaSetDevice(c->device);
	if(c->stream) cudaStreamSynchronize(c->stream);
	freeScene(c);
	if(c->stream) cudaStreamSynchronize(c->stream);
	{ cudaMemPool_t pool = nullptr; if(cudaDeviceGetDefaultMemPool(&pool, c->device) == cudaSuccess) cudaMemPoolTrimTo(pool, 0); }
	freeWavefront(c);
	cudaFree(c->d_faure); cudaFree(c->d_counters); cudaFree(c->d_fetch); cudaFree(c->d_fault); if(c->h_fault) cudaFreeHost(c->h_fault);
	for(cudaEvent_t e : c->stage_events) cudaEventDestroy(e);
	if(c->ev0) cudaEventDestroy(c->ev0);
	if(c->ev1) cudaEventDestroy(c->ev1);
	for(int k = 0; k < 2; ++k) { if(c->side[k]) { cudaStreamSynchronize(c->side[k]); cudaStreamDestroy(c->side[k]); } if(c->ev_join[k]) cudaEventDestroy(c->ev_join[k]); }
	if(c->ev_fork) cudaEventDestroy(c->ev_fork);
	if(c->stream) cudaStreamDestroy(c->stream);
	delete c;
}

void *yafgpu_stream(yafgpu_ctx *c) { return c ? (void *)c->stream : nullptr; }

int yafgpu_upload_scene(yafgpu_ctx *c, const yafgpu_scene_desc *desc)
{
	if(!c || !desc) return setErr(c, YAFGPU_ERR_INVALID, "yafgpu_upload_scene: NULL argument");
	if(desc->abi_version != YAFGPU_ABI_VERSION) return setErr(c, YAFGPU_ERR_INVALID, "yafgpu_upload_scene: ABI version mismatch");
	YG_CUDA(c, cudaSetDevice(c->device));
	YG_CUDA(c, cudaStreamSynchronize(c->stream));
	freeScene(c);
	const bool timing = getenv("YAFGPU_UPLOAD_TIMING") != nullptr;   // development aid: host wall-clock phases on stderr
	auto t_prev = std::chrono::steady_clock::now();
	auto phase = [&](const char *what) { if(!timing) return; cudaStreamSynchronize(c->stream); const auto t = std::chrono::steady_clock::now(); fprintf(stderr, "yafgpu upload: %-28s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(t - t_prev).count()); t_prev = t; };
	{	// film first: bad input must not pay for the upload and the BVH build.  Pixel coordinates travel as x | y << 16 (integrator.cu buildPixelList)
		const yafgpu_film_desc &f = desc->film;
		if(f.width <= 0 || f.height <= 0 || f.xstart < 0 || f.ystart < 0 || (long long)f.xstart + f.width > 65535 || (long long)f.ystart + f.height > 65535)
			return setErr(c, YAFGPU_ERR_INVALID, "film size: width, height > 0 and xstart + width, ystart + height <= 65535");
		if(f.premult) return setErr(c, YAFGPU_ERR_UNSUPPORTED, "premult film not implemented on the device");
	}
	const yafgpu_geometry &g = desc->geometry;
	if(g.n_tris > 0 && (!g.points || !g.tri_vidx || !g.tri_mat || !g.tri_obj || !g.objects))
		return setErr(c, YAFGPU_ERR_INVALID, "geometry arrays missing");
	if(g.n_tris > (size_t)(1u << 28)) return setErr(c, YAFGPU_ERR_UNSUPPORTED, "more than 2^28 triangles");
	// validate indices on the host (a bad index would otherwise be an illegal address on the device); 10^7 triangles are split over a few threads.  The threads run
	// WHILE this thread prepares the records and queues the host-to-device copies (6 ms of checking behind 9 ms of DMA on the bench scene); they are joined, and a bad
	// index is reported, before the first kernel reads an index (the BVH build below).
	struct IndexCheck
	{
		std::vector<std::thread> th; std::vector<const char *> bad;
		const char *join() { for(std::thread &t : th) t.join(); th.clear(); for(const char *b : bad) if(b) return b; return nullptr; }
		~IndexCheck() { for(std::thread &t : th) t.join(); }   // every early return waits for the workers: they read the caller's arrays
	} index_check;
	{
		const unsigned hw = std::thread::hardware_concurrency();
		const int n_thr = g.n_tris < (size_t)1 << 18 ? 1 : (int)std::max(1u, std::min(8u, hw ? hw : 1u));
		std::vector<const char *> &bad = index_check.bad;
		bad.assign((size_t)n_thr, nullptr);
		auto check = [&g, desc, &bad, n_thr](int t)
		{
			const size_t i0 = g.n_tris * (size_t)t / n_thr, i1 = g.n_tris * (size_t)(t + 1) / n_thr;
			for(size_t i = i0; i < i1 && !bad[t]; ++i)
			{
				for(int k = 0; k < 3; ++k)
				{
					const int v = g.tri_vidx[3 * i + k];
					if(v < 0 || (size_t)v >= g.n_points) bad[t] = "tri_vidx out of range";
					if(g.tri_nidx) { const int q = g.tri_nidx[3 * i + k]; if(q >= (int)g.n_normals) bad[t] = "tri_nidx out of range"; }
					if(g.tri_uvidx) { const int q = g.tri_uvidx[3 * i + k]; if(q >= (int)g.n_uv) bad[t] = "tri_uvidx out of range"; }
				}
				if(g.tri_mat[i] < 0 || (size_t)g.tri_mat[i] >= desc->n_materials) bad[t] = "tri_mat out of range";
				if(g.tri_obj[i] < 0 || (size_t)g.tri_obj[i] >= g.n_objects) bad[t] = "tri_obj out of range";
			}
		};
		if(n_thr == 1) { check(0); if(bad[0]) return setErr(c, YAFGPU_ERR_INVALID, bad[0]); }
		else for(int t = 0; t < n_thr; ++t) index_check.th.emplace_back(check, t);
	}

	phase("index validation (started)");
	// materials / lights / camera / film: reference constructors restated on the host (host_prep.h)
	c->h_mats.assign(desc->n_materials, MatDev());
	std::vector<unsigned> triflags(desc->n_materials ? desc->n_materials : 1, 0u);
	std::vector<TexDev> h_tex(desc->n_textures);
	for(size_t i = 0; i < desc->n_textures; ++i)
	{
		std::string err;
		if(!desc->textures || !prepTexture(desc->textures[i], h_tex[i], err)) return setErr(c, YAFGPU_ERR_UNSUPPORTED, "texture " + std::to_string(i) + ": " + err);
	}
	std::vector<NodeDev> h_nodes;
	for(size_t i = 0; i < desc->n_materials; ++i)
	{
		std::string err;
		if(!prepNodes(desc->materials[i], *desc, h_tex, c->h_mats[i], h_nodes, err)) return setErr(c, YAFGPU_ERR_UNSUPPORTED, "material " + std::to_string(i) + ": " + err);
		if(!prepMaterial(desc->materials[i], c->h_mats[i], err)) return setErr(c, YAFGPU_ERR_UNSUPPORTED, "material " + std::to_string(i) + ": " + err);
	}
	for(size_t i = 0; i < desc->n_materials; ++i)
	{
		std::string err;
		if((c->h_mats[i].type == YAFGPU_MAT_BLEND || c->h_mats[i].type == YAFGPU_MAT_MASK) && !finishBlend(c->h_mats, i, err)) return setErr(c, YAFGPU_ERR_UNSUPPORTED, "material " + std::to_string(i) + ": " + err);
	}
	for(size_t i = 0; i < desc->n_materials; ++i) { c->h_mats[i].obj_flags = c->h_mats[i].bsdf_flags; c->h_mats[i].is_mask = 0; }
	bool any_dispersion = false;
	for(size_t i = 0; i < desc->n_materials; ++i)
	{
		const MatDev &md = c->h_mats[i];
		if(!md.disperse) continue;
		any_dispersion = true;
		// with an IOR shader the reference recomputes the Cauchy coefficients per hit (material_glass.cc:92-95)
		if(md.node_count > 0 && md.slot[YAFGPU_SLOT_IOR] >= 0) return setErr(c, YAFGPU_ERR_UNSUPPORTED, "material " + std::to_string(i) + ": dispersion together with an IOR shader is not implemented on the device");
	}
	for(size_t i = 0; i < desc->n_materials; ++i)
		if((c->h_mats[i].type == YAFGPU_MAT_BLEND || c->h_mats[i].type == YAFGPU_MAT_MASK) && (c->h_mats[c->h_mats[i].mat1].disperse || c->h_mats[c->h_mats[i].mat2].disperse))
			return setErr(c, YAFGPU_ERR_UNSUPPORTED, "material " + std::to_string(i) + ": blend / mask over a dispersive glass is not implemented on the device");
	if(any_dispersion && (!desc->cie_xyz || desc->n_cie_rows < 471)) return setErr(c, YAFGPU_ERR_INVALID, "a material has dispersion_power > 0 but yafgpu_scene_desc.cie_xyz (471 rows) is missing");
	for(size_t i = 0; i < desc->n_materials; ++i) triflags[i] = triFlagsOf(c->h_mats[i]);
	for(size_t i = 0; i < desc->n_materials; ++i)
	{	// Scene::isShadowed calls Material::getTransparency without initBsdf (kdtree_triangle.cc:1106-1124): with shader nodes it would read
		// the node stack of whatever material was shaded last.  Not reproducible, so the combination is refused.
		const MatDev &md = c->h_mats[i];
		const bool transp_nodes = md.node_count > 0 && md.is_transparent;
		const bool blend_child_nodes = md.type == YAFGPU_MAT_BLEND && (c->h_mats[md.mat1].node_count > 0 || c->h_mats[md.mat2].node_count > 0);
		if(blend_child_nodes) return setErr(c, YAFGPU_ERR_UNSUPPORTED, "material " + std::to_string(i) + ": blend_mat of materials with shader nodes is not implemented on the device");
		if(transp_nodes && desc->integrator.transp_shad) return setErr(c, YAFGPU_ERR_UNSUPPORTED, "material " + std::to_string(i) + ": transparent shadows through a material with shader nodes");
	}
	c->h_lights.clear();
	std::vector<size_t> light_src; // descriptor index of every kept light; disabled lights are never added (Scene::addLight scene.cc:688-698)
	for(size_t i = 0; i < desc->n_lights; ++i)
	{
		if(!desc->lights[i].light_enabled) continue;
		std::string err;
		LightDev ld;
		if(!prepLight(desc->lights[i], ld, err)) return setErr(c, YAFGPU_ERR_UNSUPPORTED, "light " + std::to_string(i) + ": " + err);
		if(ld.object >= (int)g.n_objects) return setErr(c, YAFGPU_ERR_INVALID, "light object index out of range");
		if((ld.type == YAFGPU_LIGHT_MESH || ld.type == YAFGPU_LIGHT_BGPORTAL) && ld.object < 0) return setErr(c, YAFGPU_ERR_INVALID, "meshlight / bgPortalLight without a valid object");
		if(ld.type == YAFGPU_LIGHT_BGPORTAL && desc->background.type == YAFGPU_BG_NONE) return setErr(c, YAFGPU_ERR_INVALID, "bgPortalLight without a background");
		c->h_lights.push_back(ld); light_src.push_back(i);
	}
	SceneDev &s = c->scene;
	memset(&s, 0, sizeof(s));
	{
		std::string err;
		if(!prepCamera(desc->camera, s.cam, err)) return setErr(c, YAFGPU_ERR_UNSUPPORTED, err);
	}
	const yafgpu_integrator_desc &in = desc->integrator;
	if(in.type != YAFGPU_INT_DIRECTLIGHT && in.type != YAFGPU_INT_PATHTRACING) return setErr(c, YAFGPU_ERR_UNSUPPORTED, "integrator type");
	// bg_transp: the alpha of a camera sample is 0 when its primary ray leaves the scene, 1 otherwise (integrator_path_tracer.cc:122-123,324;
	// integrator_direct_light.cc likewise).  bg_transp_refract folds Material::getAlpha along the refraction chain (YG_ST_ALPHA in integrator.cu).
	// the packed task state holds the path depth in 6 bits (integrator.cu packState); the recursion level never passes 19 — the reference's own
	// `raylevel_ < 20` guards (integrator_montecarlo.cc:862,921) — so the wavefront needs at most 20 + bounces + 1 iterations
	if(in.bounces > 63) return setErr(c, YAFGPU_ERR_UNSUPPORTED, "bounces > 63: the wavefront task state holds the path depth in 6 bits");
	if(in.raydepth < 0 || in.bounces < 0) return setErr(c, YAFGPU_ERR_INVALID, "negative raydepth / bounces");
	if(std::min(in.raydepth + 31, 19) + 1 + in.bounces + 1 > YG_MAX_ITERS) return setErr(c, YAFGPU_ERR_UNSUPPORTED, "raydepth + bounces too large for the wavefront schedule");
	s.integ.type = in.type; s.integ.r_depth = in.raydepth; s.integ.tr_shad = in.transp_shad; s.integ.s_depth = in.shadow_depth;
	s.integ.n_paths = in.path_samples; s.integ.max_bounces = in.bounces; s.integ.rr_min_bounces = in.rr_min_bounces;
	s.integ.no_recursive = in.no_recursive; s.integ.trace_caustics = in.trace_caustics;
	s.integ.transp_bg = in.bg_transp ? 1 : 0; s.integ.transp_refr_bg = in.bg_transp_refract ? 1 : 0;
	s.integ.do_ao = in.do_ao; s.integ.ao_samples = in.ao_samples; s.integ.ao_dist = in.ao_distance;
	for(int k = 0; k < 3; ++k) s.integ.ao_col[k] = in.ao_color[k];
	if(in.do_ao && in.ao_samples < 1) return setErr(c, YAFGPU_ERR_INVALID, "AO_samples must be positive");
	s.integ.shadow_bias = in.shadow_bias; s.integ.shadow_bias_auto = in.shadow_bias_auto;
	s.integ.ray_min_dist = in.ray_min_dist; s.integ.ray_min_dist_auto = in.ray_min_dist_auto;
	s.integ.aa_light_sample_multiplier = desc->film.aa_light_sample_multiplier;
	s.integ.aa_indirect_sample_multiplier = desc->film.aa_indirect_sample_multiplier;
	s.integ.aa_clamp_indirect = desc->film.aa_clamp_indirect;
	// Scene::update scene.cc:823-824
	if(in.shadow_bias_auto) s.integ.shadow_bias = (float)0.0005;
	if(in.ray_min_dist_auto) s.integ.ray_min_dist = (float)0.00005;

	const yafgpu_background &bg = desc->background;
	s.bg.type = bg.type;
	if(bg.type == YAFGPU_BG_CONSTANT) for(int k = 0; k < 3; ++k) s.bg.color[k] = bg.color[k] * bg.power; // background_constant.cc:86
	else if(bg.type == YAFGPU_BG_TEXTURE)
	{	// TextureBackground ctor background_texture.cc:32-39
		if(!bg.tex_rgb || bg.tex_w <= 0 || bg.tex_h <= 0) return setErr(c, YAFGPU_ERR_INVALID, "texture background without texels");
		s.bg.mapping_angular = bg.mapping_angular; s.bg.power = bg.power; s.bg.tex_interp = bg.tex_interp;
		s.bg.rotation = 2.0f * bg.rotation / 360.f;
		s.bg.sin_r = fSin((float)(YG_M_PI * (double)s.bg.rotation)); s.bg.cos_r = fCos((float)(YG_M_PI * (double)s.bg.rotation));
		s.bg.tex_w = bg.tex_w; s.bg.tex_h = bg.tex_h;
		s.bg.tex = bg.tex_rgb; // host pointer while the light tables are built below; replaced by the device copy afterwards
	}
	else if(bg.type == YAFGPU_BG_GRADIENT)
	{	// GradientBackground::factory background_gradient.cc:88: every colour * power; the constructor is handed with_caustic = true
		for(int k = 0; k < 3; ++k)
		{
			s.bg.gzenith[k] = bg.zenith_ground_color[k] * bg.power; s.bg.ghoriz[k] = bg.horizon_ground_color[k] * bg.power;
			s.bg.szenith[k] = bg.zenith_color[k] * bg.power; s.bg.shoriz[k] = bg.horizon_color[k] * bg.power;
		}
	}
	else if(bg.type != YAFGPU_BG_NONE) return setErr(c, YAFGPU_ERR_UNSUPPORTED, "background type not implemented on the device yet");
	s.bg.has_ibl = bg.ibl; s.bg.shoots_caustic = bg.type == YAFGPU_BG_GRADIENT ? 1 : bg.with_caustic;
	// light tables: MeshLight::initIs area CDFs, BackgroundLight::init row/column CDFs
	std::vector<float> light_tab;
	for(const LightDev &ld : c->h_lights)
		if(ld.type == YAFGPU_LIGHT_MESH || ld.type == YAFGPU_LIGHT_BGPORTAL)
		{	// the area tables of a mesh light are built on the host from the index arrays: they must have been checked by now
			if(const char *b = index_check.join()) return setErr(c, YAFGPU_ERR_INVALID, b);
			break;
		}
	for(size_t i = 0; i < c->h_lights.size(); ++i)
	{
		LightDev &ld = c->h_lights[i];
		if(ld.type == YAFGPU_LIGHT_MESH || ld.type == YAFGPU_LIGHT_BGPORTAL) buildMeshLightTables(g, g.objects[ld.object], ld, light_tab);
		else if(ld.type == YAFGPU_LIGHT_IES)
		{
			std::string ies_err;
			if(!buildIesTables(desc->lights[i], ld, light_tab, ies_err)) return setErr(c, YAFGPU_ERR_INVALID, ies_err.c_str());
		}
		else if(ld.type == YAFGPU_LIGHT_BACKGROUND)
		{
			if(bg.type == YAFGPU_BG_NONE) return setErr(c, YAFGPU_ERR_INVALID, "bglight without a background");
			buildBgLightTables(s.bg, ld, light_tab);
		}
	}

	// tag objects with their light (AreaLight::init light_area.cc:55-63): the LAST light naming the object wins, as setLight overwrites
	std::vector<ObjDev> objs(g.n_objects ? g.n_objects : 1);
	for(size_t i = 0; i < g.n_objects; ++i) { objs[i].has_orco = g.objects[i].has_orco; objs[i].has_uv = g.objects[i].has_uv; objs[i].smooth = g.objects[i].smooth; objs[i].light = g.objects[i].light; }
	for(size_t i = 0; i < c->h_lights.size(); ++i)
		if((c->h_lights[i].type == YAFGPU_LIGHT_AREA || c->h_lights[i].type == YAFGPU_LIGHT_MESH || c->h_lights[i].type == YAFGPU_LIGHT_BGPORTAL) && c->h_lights[i].object >= 0) objs[c->h_lights[i].object].light = (int)i;

	cudaStream_t st = c->stream;
	c->n_tris = g.n_tris; c->n_points = g.n_points; c->n_normals = g.n_normals; c->n_uv = g.n_uv; c->n_objects = g.n_objects;
	c->n_mats = desc->n_materials; c->n_lights = c->h_lights.size();
	struct TmpIdx   // index arrays only k_tri_shade needs: released on every way out of this function
	{
		int *nidx = nullptr, *uvidx = nullptr, *obj = nullptr; cudaStream_t s;
		~TmpIdx() { poolFree(nidx, s); poolFree(uvidx, s); poolFree(obj, s); }
	} tmp;
	tmp.s = st;
	int *&d_nidx = tmp.nidx, *&d_uvidx = tmp.uvidx, *&d_obj = tmp.obj;
	YG_CUDA(c, upload(c->d_points, g.points, g.n_points * 3, st));
	YG_CUDA(c, upload(c->d_normals, g.normals, g.n_normals * 3, st));
	YG_CUDA(c, upload(c->d_uv, g.uv, g.n_uv * 2, st));
	YG_CUDA(c, upload(c->d_tri_vidx, g.tri_vidx, g.n_tris * 3, st));
	YG_CUDA(c, upload(c->d_tri_mat, g.tri_mat, g.n_tris, st));
	YG_CUDA(c, upload(c->d_mat_triflags, triflags.data(), triflags.size(), st));
	{	// sort buckets of the shading kernel: materials ranked by (type, index) so that equal code paths are adjacent
		std::vector<int> order(c->h_mats.size()), bucket(c->h_mats.size() ? c->h_mats.size() : 1, 0);
		for(size_t i = 0; i < order.size(); ++i) order[i] = (int)i;
		auto ktype = [&](int i) { return c->h_mats[i].node_count > 0 ? YG_MAT_NODES : c->h_mats[i].type; }; // kernel that shades the material
		std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return ktype(a) < ktype(b); });
		// one bucket per material while they fit, else one per type: a bucket never mixes types (per-type shading kernels)
		const bool per_mat = order.size() <= 4096;
		c->n_mat_buckets = per_mat ? (int)std::max<size_t>(order.size(), 1) : YG_N_MAT_TYPES;
		for(int t = 0; t < YG_N_MAT_TYPES; ++t) { c->type_b0[t] = 0; c->type_b1[t] = -1; }
		for(size_t r = 0; r < order.size(); ++r)
		{
			const int t = ktype(order[r]), b = per_mat ? (int)r : t;
			bucket[order[r]] = b;
			if(c->type_b1[t] < 0) c->type_b0[t] = b;
			c->type_b1[t] = b;
		}
		YG_CUDA(c, upload(c->d_mat_bucket, bucket.data(), bucket.size(), st));
	}
	YG_CUDA(c, upload(c->d_objects, objs.data(), objs.size(), st));
	YG_CUDA(c, upload(c->d_mats, c->h_mats.data(), c->h_mats.size(), st));
	bool any_noise = false;
	for(const TexDev &t : h_tex) any_noise = any_noise || t.type >= YAFGPU_TEX_CLOUDS;
	if(any_noise)
	{
		if(!desc->noise_tables || desc->noise_tables_bytes != YAFGPU_NOISE_TABLES_BYTES)
			return setErr(c, YAFGPU_ERR_INVALID, "a noise-based procedural texture is present but yafgpu_scene_desc.noise_tables (13340 bytes) is missing");
		YG_CUDA(c, upload(c->d_noise, (const NoiseTables *)desc->noise_tables, 1, st));
	}
	for(size_t i = 0; i < h_tex.size(); ++i)
	{
		float4 *d_px = nullptr;
		h_tex[i].noise = c->d_noise;
		h_tex[i].ramp = nullptr;
		if(h_tex[i].n_ramp > 0)
		{
			float *d_ramp = nullptr;
			YG_CUDA(c, upload(d_ramp, desc->textures[i].ramp, (size_t)5 * h_tex[i].n_ramp, st));
			c->tex_ptrs.push_back(d_ramp);
			h_tex[i].ramp = d_ramp;
		}
		if(h_tex[i].type != YAFGPU_TEX_IMAGE) { h_tex[i].tex = nullptr; continue; } // procedural: no texels
		YG_CUDA(c, upload(d_px, (const float4 *)desc->textures[i].rgba, (size_t)h_tex[i].w * h_tex[i].h, st));
		c->tex_ptrs.push_back(d_px);
		h_tex[i].tex = d_px;
	}
	YG_CUDA(c, upload(c->d_textures, h_tex.data(), h_tex.size(), st));
	YG_CUDA(c, upload(c->d_shader_nodes, h_nodes.data(), h_nodes.size(), st));
	if(any_dispersion) YG_CUDA(c, upload(c->d_cie, desc->cie_xyz, (size_t)471 * 3, st));
	YG_CUDA(c, poolAlloc(&c->d_tri_ng, sizeof(float) * 3 * (g.n_tris ? g.n_tris : 1), st));
	YG_CUDA(c, poolAlloc(&c->d_tris, sizeof(TriShade) * (g.n_tris ? g.n_tris : 1), st));
	if(g.tri_nidx) YG_CUDA(c, upload(d_nidx, g.tri_nidx, g.n_tris * 3, st));
	if(g.tri_uvidx) YG_CUDA(c, upload(d_uvidx, g.tri_uvidx, g.n_tris * 3, st));
	YG_CUDA(c, upload(d_obj, g.tri_obj, g.n_tris, st));

	if(const char *b = index_check.join()) return setErr(c, YAFGPU_ERR_INVALID, b);
	phase("host prep + H2D copies + index validation");
	// K1: accelerator
	cudaEventRecord(c->ev0, st);
	Accel &a = s.accel;
	a.n_tris = (int)g.n_tris;
	a.fault = c->d_fault; a.stack_limit = c->stack_limit > 0 ? c->stack_limit : YG_STACK;
	if(g.n_tris > 0)
	{
		BvhBuildInput bi; bi.n_tris = (int)g.n_tris; bi.points = c->d_points; bi.tri_vidx = c->d_tri_vidx; bi.tri_mat = c->d_tri_mat;
		bi.mat_triflags = c->d_mat_triflags; bi.tri_ng_out = c->d_tri_ng; bi.leaf_max = c->leaf_max; bi.width = YG_BVH_WIDTH;
		BvhBuildOutput bo;
		cudaError_t e = buildBvh(bi, bo, st);
		if(e != cudaSuccess) return cudaErr(c, e, "buildBvh");
		c->d_nodes = bo.nodes; c->d_acc_tris = bo.tris;
		a.nodes = (const uint4 *)bo.nodes; a.tris = bo.tris; a.n_nodes = bo.n_nodes; a.root = bo.root; a.width = YG_BVH_WIDTH;
		c->stats.bvh_node_bytes = (uint32_t)bo.node_bytes; c->stats.bvh_width = YG_BVH_WIDTH; c->stats.bvh_levels = (uint32_t)bo.levels;
		memcpy(c->scene_bound, bo.bound, sizeof(bo.bound));
		// tree bound grown by 0.1 % per axis (kdtree_triangle.cc:112-116; double arithmetic on float members)
		for(int k = 0; k < 3; ++k)
		{
			double foo = (double)(bo.bound[3 + k] - bo.bound[k]) * 0.001;
			a.bmin[k] = (float)((double)bo.bound[k] - foo); a.bmax[k] = (float)((double)bo.bound[3 + k] + foo);
			c->scene_bound[k] = a.bmin[k]; c->scene_bound[3 + k] = a.bmax[k]; // Scene::scene_bound_ = tree_->getBound() scene.cc:820
		}
		k_tri_shade<<<(unsigned)((g.n_tris + 255) / 256), 256, 0, st>>>((int)g.n_tris, c->d_tri_vidx, d_nidx, d_uvidx, c->d_tri_mat, d_obj, c->d_tri_ng, c->d_tris);
	}
	phase("BVH build");
	// mesh lights intersect their own accelerator (MeshLight::initIs builds a private TriKdTree, light_meshlight.cc:72)
	std::vector<Accel> laccels;
	for(size_t i = 0; i < c->h_lights.size(); ++i)
	{
		LightDev &ld = c->h_lights[i];
		if((ld.type != YAFGPU_LIGHT_MESH && ld.type != YAFGPU_LIGHT_BGPORTAL) || ld.tri_count <= 0) continue;
		BvhBuildInput bi; bi.n_tris = ld.tri_count; bi.points = c->d_points; bi.tri_vidx = c->d_tri_vidx + 3 * (size_t)ld.tri_first;
		bi.tri_mat = c->d_tri_mat + ld.tri_first; bi.mat_triflags = c->d_mat_triflags; bi.tri_ng_out = c->d_tri_ng + 3 * (size_t)ld.tri_first; bi.leaf_max = c->leaf_max; bi.width = YG_BVH_WIDTH;
		BvhBuildOutput bo;
		cudaError_t e2 = buildBvh(bi, bo, st);
		if(e2 != cudaSuccess) return cudaErr(c, e2, "buildBvh (mesh light)");
		Accel la; memset(&la, 0, sizeof(la));
		la.nodes = (const uint4 *)bo.nodes; la.tris = bo.tris; la.n_nodes = bo.n_nodes; la.n_tris = ld.tri_count; la.root = bo.root; la.width = YG_BVH_WIDTH; la.fault = c->d_fault; la.stack_limit = a.stack_limit;
		c->light_bvh_ptrs.push_back(bo.nodes); c->light_bvh_ptrs.push_back(bo.tris);
		ld.accel = (int)laccels.size();
		laccels.push_back(la);
	}
	YG_CUDA(c, upload(c->d_light_accels, laccels.data(), laccels.size(), st));
	YG_CUDA(c, upload(c->d_lights, c->h_lights.data(), c->h_lights.size(), st));
	YG_CUDA(c, upload(c->d_light_tables, light_tab.data(), light_tab.size(), st));
	if(bg.type == YAFGPU_BG_TEXTURE)
	{
		YG_CUDA(c, upload(c->d_bg_tex, bg.tex_rgb, (size_t)3 * bg.tex_w * bg.tex_h, st));
		s.bg.tex = c->d_bg_tex;
	}
	cudaEventRecord(c->ev1, st);
	cudaError_t e = cudaStreamSynchronize(st);
	if(e != cudaSuccess) return cudaErr(c, e, "scene upload");
	float ms = 0.f; cudaEventElapsedTime(&ms, c->ev0, c->ev1); c->stats.ms_build += ms;
	c->stats.bvh_nodes = a.n_nodes; c->stats.bvh_bytes = (uint64_t)a.n_nodes * c->stats.bvh_node_bytes; c->stats.tri_bytes = (uint64_t)g.n_tris * sizeof(TriAccel);

	s.points = c->d_points; s.normals = c->d_normals; s.uv = c->d_uv; s.tris = c->d_tris; s.objects = c->d_objects;
	s.materials = c->d_mats; s.lights = c->d_lights; s.n_lights = (int)c->h_lights.size(); s.light_tables = c->d_light_tables; s.light_accels = c->d_light_accels; s.faure = c->d_faure;
	s.textures = c->d_textures; s.nodes = c->d_shader_nodes; s.cie_xyz = c->d_cie; s.n_cie = c->d_cie ? 471 : 0;

	// film (ImageFilm ctor + init)
	prepFilm(desc->film, c->film);
	c->film.transp_bg = s.integ.transp_bg;
	c->tile_size = desc->film.tile_size > 0 ? desc->film.tile_size : 32;
	YG_CUDA(c, poolAlloc(&c->film.pix, sizeof(float) * 5 * (size_t)c->film.w * c->film.h, st));
	YG_CUDA(c, cudaMemsetAsync(c->film.pix, 0, sizeof(float) * 5 * (size_t)c->film.w * c->film.h, st));
	YG_CUDA(c, cudaStreamSynchronize(st));
	phase("lights, film, rest");
	c->wave_cur_cap = 0; c->iters_hint = 0; c->aa_mask_valid = false; c->aa_mask.clear();
	c->uploaded = true;
	return YAFGPU_OK;
}

int yafgpu_scene_bound(yafgpu_ctx *c, float *b)
{
	if(!c || !b) return setErr(c, YAFGPU_ERR_INVALID, "NULL argument");
	if(!c->uploaded) return setErr(c, YAFGPU_ERR_STATE, "no scene uploaded");
	memcpy(b, c->scene_bound, sizeof(float) * 6);
	return YAFGPU_OK;
}

int yafgpu_trace_device(yafgpu_ctx *c, const void *d_rays, size_t n, void *d_hits, int mode, int shadow_depth)
{
	if(!c) return YAFGPU_ERR_INVALID;
	if(!c->uploaded) return setErr(c, YAFGPU_ERR_STATE, "no scene uploaded");
	if(mode == YAFGPU_TRACE_SHADOW_TRANSP) return setErr(c, YAFGPU_ERR_UNSUPPORTED, "transparent-shadow batch trace is only reachable through the integrator");
	(void)shadow_depth;
	if(n == 0) return YAFGPU_OK;
	YG_CUDA(c, cudaSetDevice(c->device));
	int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
	if(n > 0x7fffffffu) return setErr(c, YAFGPU_ERR_INVALID, "at most 2^31 - 1 rays per call");
	YG_CUDA(c, cudaMemsetAsync(c->d_fetch, 0, sizeof(int), c->stream));
	size_t want = (n + 127) / 128;
	unsigned grid = (unsigned)(want < (size_t)sms * 16 ? want : (size_t)sms * 16);
	if(c->flags & YAFGPU_FLAG_COUNT_TRAVERSAL)
		k_trace_batch<true><<<grid, 128, 0, c->stream>>>(c->scene.accel, (const yafgpu_ray *)d_rays, n, (yafgpu_hit *)d_hits, mode, c->d_fetch, c->regen_threshold, c->d_counters);
	else
		k_trace_batch<false><<<grid, 128, 0, c->stream>>>(c->scene.accel, (const yafgpu_ray *)d_rays, n, (yafgpu_hit *)d_hits, mode, c->d_fetch, c->regen_threshold, c->d_counters);
	YG_CUDA(c, cudaGetLastError());
	c->stats.kernel_launches++;
	if(mode == YAFGPU_TRACE_CLOSEST) c->stats.rays_extension += n; else c->stats.rays_shadow += n;
	return YAFGPU_OK;
}

int yafgpu_trace(yafgpu_ctx *c, const yafgpu_ray *rays, size_t n, yafgpu_hit *hits, int mode, int shadow_depth)
{
	if(!c || (n && (!rays || !hits))) return setErr(c, YAFGPU_ERR_INVALID, "NULL argument");
	if(!c->uploaded) return setErr(c, YAFGPU_ERR_STATE, "no scene uploaded");
	if(n == 0) return YAFGPU_OK;
	YG_CUDA(c, cudaSetDevice(c->device));
	void *d_r = nullptr, *d_h = nullptr;
	YG_CUDA(c, poolAlloc(&d_r, n * sizeof(yafgpu_ray), c->stream));
	cudaError_t e = poolAlloc(&d_h, n * sizeof(yafgpu_hit), c->stream);
	if(e != cudaSuccess) { poolFree(d_r, c->stream); return cudaErr(c, e, "device allocation for hits"); }
	e = cudaMemcpyAsync(d_r, rays, n * sizeof(yafgpu_ray), cudaMemcpyHostToDevice, c->stream);
	int rc = (e == cudaSuccess) ? yafgpu_trace_device(c, d_r, n, d_h, mode, shadow_depth) : cudaErr(c, e, "H2D rays");
	if(rc == YAFGPU_OK)
	{
		e = cudaMemcpyAsync(hits, d_h, n * sizeof(yafgpu_hit), cudaMemcpyDeviceToHost, c->stream);
		if(e == cudaSuccess) e = cudaMemcpyAsync(c->h_fault, c->d_fault, sizeof(int), cudaMemcpyDeviceToHost, c->stream);
		if(e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
		if(e != cudaSuccess) rc = cudaErr(c, e, "D2H hits");
		else rc = faultCheck(c, false);
	}
	poolFree(d_r, c->stream); poolFree(d_h, c->stream);
	return rc;
}

// Device address of the static scratch array on `device` (which must be current): looked up once per device.
static std::mutex g_static_mutex[YG_MAX_DEVICES];
static unsigned char *g_static_ptr[YG_MAX_DEVICES];
static cudaError_t staticScratch(int device, unsigned char **out)
{
	if(!g_static_ptr[device])
	{
		void *p = nullptr;
		cudaError_t e = cudaGetSymbolAddress(&p, g_trace_static_scratch);
		if(e != cudaSuccess) return e;
		g_static_ptr[device] = (unsigned char *)p;
	}
	*out = g_static_ptr[device];
	return cudaSuccess;
}

int yafgpu_trace_static(yafgpu_ctx *c, const yafgpu_ray *rays, size_t n, yafgpu_hit *hits, int mode)
{
	if(!c || (n && (!rays || !hits))) return setErr(c, YAFGPU_ERR_INVALID, "NULL argument");
	if(!c->uploaded) return setErr(c, YAFGPU_ERR_STATE, "no scene uploaded");
	if(mode != YAFGPU_TRACE_CLOSEST && mode != YAFGPU_TRACE_SHADOW) return setErr(c, YAFGPU_ERR_UNSUPPORTED, "yafgpu_trace_static: mode must be YAFGPU_TRACE_CLOSEST or YAFGPU_TRACE_SHADOW");
	if(n > (size_t)YAFGPU_TRACE_STATIC_MAX_RAYS)
		return setErr(c, YAFGPU_ERR_INVALID, "yafgpu_trace_static: " + std::to_string(n) + " rays exceed the static scratch capacity of " + std::to_string(YAFGPU_TRACE_STATIC_MAX_RAYS) +
		              " rays per call (YAFGPU_TRACE_STATIC_MAX_RAYS); split the batch or use yafgpu_trace");
	if(n == 0) return YAFGPU_OK;
	if(c->device < 0 || c->device >= YG_MAX_DEVICES) return setErr(c, YAFGPU_ERR_INVALID, "yafgpu_trace_static: device index beyond the per-device table");
	YG_CUDA(c, cudaSetDevice(c->device));
	// every call on this device shares the one array: callers of this process are serialised here, through the final synchronise
	std::lock_guard<std::mutex> lock(g_static_mutex[c->device]);
	unsigned char *scratch = nullptr;
	YG_CUDA(c, staticScratch(c->device, &scratch));
	unsigned char *d_rays = scratch + 256, *d_hits = scratch + 256 + (size_t)YAFGPU_TRACE_STATIC_MAX_RAYS * sizeof(yafgpu_ray);
	// the array keeps the previous call's contents: the fetch counter must start at zero; rays [0, n) are overwritten and every hit
	// record in [0, n) is written by the kernel, nothing beyond n is read
	YG_CUDA(c, cudaMemsetAsync(scratch, 0, 256, c->stream));
	YG_CUDA(c, cudaMemcpyAsync(d_rays, rays, n * sizeof(yafgpu_ray), cudaMemcpyHostToDevice, c->stream));
	int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
	const size_t want = (n + 127) / 128;
	const unsigned grid = (unsigned)(want < (size_t)sms * 16 ? want : (size_t)sms * 16);
	k_trace_batch_static<<<grid, 128, 0, c->stream>>>(c->scene.accel, scratch, (int)n, mode, c->regen_threshold);
	YG_CUDA(c, cudaGetLastError());
	YG_CUDA(c, cudaMemcpyAsync(hits, d_hits, n * sizeof(yafgpu_hit), cudaMemcpyDeviceToHost, c->stream));
	YG_CUDA(c, cudaMemcpyAsync(c->h_fault, c->d_fault, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
	YG_CUDA(c, cudaMemcpyAsync(c->h_fault + 1, scratch, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
	YG_CUDA(c, cudaStreamSynchronize(c->stream));
	c->stats.kernel_launches++;
	if(mode == YAFGPU_TRACE_CLOSEST) c->stats.rays_extension += n; else c->stats.rays_shadow += n;
	// every warp reserves rays until the counter passes n: a final value below n means somebody else reset it under the kernel
	if((size_t)c->h_fault[1] < n) return setErr(c, YAFGPU_ERR_CUDA, "yafgpu_trace_static: the static scratch was disturbed while the call ran (fetch counter " + std::to_string(c->h_fault[1]) + " < " + std::to_string(n) + " rays): overlapping use of the array");
	return faultCheck(c, false);
}

int yafgpu_film_clear(yafgpu_ctx *c)
{
	if(!c || !c->uploaded) return setErr(c, YAFGPU_ERR_STATE, "no scene uploaded");
	YG_CUDA(c, cudaSetDevice(c->device));
	YG_CUDA(c, cudaMemsetAsync(c->film.pix, 0, sizeof(float) * 5 * (size_t)c->film.w * c->film.h, c->stream));
	return YAFGPU_OK;
}

int yafgpu_render_pass(yafgpu_ctx *c, const yafgpu_pass_desc *pass)
{
	if(!c || !pass) return setErr(c, YAFGPU_ERR_INVALID, "NULL argument");
	if(!c->uploaded) return setErr(c, YAFGPU_ERR_STATE, "no scene uploaded");
	YG_CUDA(c, cudaSetDevice(c->device));
	if(pass->adaptive == 2)
	{	// the mask of the last yafgpu_next_pass; an empty one stands for "every pixel" (threshold <= 0)
		if(!c->aa_mask_valid) return setErr(c, YAFGPU_ERR_STATE, "adaptive = 2 needs a yafgpu_next_pass on this scene first");
		yafgpu_pass_desc p = *pass;
		p.adaptive = 1; p.resample_mask = c->aa_mask.empty() ? nullptr : c->aa_mask.data();
		return renderPass(c, &p);
	}
	return renderPass(c, pass);
}

} // extern "C"

namespace yg {
// ---- ImageFilm::nextPass on the device (imagefilm.cc:306-424) ---------------------------------------------------------------------------------
struct AaPix { float r, g, b, a; };
// Pixel::normalized util_image_buffers.h:39-43 with Rgba / float of color.h:310-314 (the reciprocal is formed in double and narrowed)
__device__ __forceinline__ AaPix aaNormalized(const float *__restrict__ film, int w, int x, int y)
{
	const float *p = film + 5 * ((size_t)y * w + x);
	float f = p[4];
	if(f == 0.f) return AaPix{0.f, 0.f, 0.f, 0.f};
	f = (float)(1.0 / (double)f);
	return AaPix{p[0] * f, p[1] * f, p[2] * f, p[3] * f};
}
__device__ __forceinline__ float aaBri(const AaPix &c) { return 0.2126f * c.r + 0.7152f * c.g + 0.0722f * c.b; }   // Rgb::col2Bri color.h:60
// Rgba::colorDifference color.h:447-465
__device__ __forceinline__ float aaDifference(const AaPix &a, const AaPix &b, int rgb_too)
{
	float d = fabsf(aaBri(b) - aaBri(a));
	if(rgb_too)
	{
		const float rd = fabsf(b.r - a.r), gd = fabsf(b.g - a.g), bd = fabsf(b.b - a.b), ad = fabsf(b.a - a.a);
		if(d < rd) d = rd;
		if(d < gd) d = gd;
		if(d < bd) d = bd;
		if(d < ad) d = ad;
	}
	return d;
}
// ImageFilm::darkThresholdCurveInterpolate imagefilm.cc:1312-1328: a piecewise-linear threshold over the pixel brightness
__device__ __forceinline__ float aaDarkCurve(float b)
{
	const float kx[15] = {0.10f, 0.20f, 0.30f, 0.40f, 0.50f, 0.60f, 0.70f, 0.80f, 0.90f, 1.00f, 1.20f, 1.40f, 1.80f, 0.f, 0.f};
	const float ky[15] = {0.0001f, 0.0010f, 0.0020f, 0.0035f, 0.0055f, 0.0075f, 0.0100f, 0.0150f, 0.0250f, 0.0400f, 0.0800f, 0.0950f, 0.1000f, 0.f, 0.f};
	const float kw[15] = {0.10f, 0.10f, 0.10f, 0.10f, 0.10f, 0.10f, 0.10f, 0.10f, 0.10f, 0.20f, 0.20f, 0.40f, 0.f, 0.f, 0.f};
	if(b <= 0.10f) return 0.0001f;
	for(int k = 0; k < 12; ++k)
		if(b > kx[k] && b <= kx[k + 1]) return (ky[k] + (b - kx[k]) * (ky[k + 1] - ky[k]) / kw[k]);
	return 0.1000f;
}
__global__ void __launch_bounds__(256) k_next_pass(const float *__restrict__ film, int w, int h, yafgpu_aa_noise_params prm, unsigned char *__restrict__ flags)
{
	const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
	if(x >= w - 1 || y >= h - 1) return;
	if(film[5 * ((size_t)y * w + x) + 4] <= 0.f) flags[(size_t)y * w + x] = 1;   // never rendered (a film loaded from disk): sample it
	const AaPix c = aaNormalized(film, w, x, y);
	const float bri = 0.2126f * fabsf(c.r) + 0.7152f * fabsf(c.g) + 0.0722f * fabsf(c.b);   // abscol2Bri
	float thr = prm.threshold;
	if(prm.dark_detection_type == YAFGPU_DARK_LINEAR && prm.dark_threshold_factor > 0.f) thr = prm.threshold * ((1.f - prm.dark_threshold_factor) + (bri * prm.dark_threshold_factor));
	else if(prm.dark_detection_type == YAFGPU_DARK_CURVE) thr = aaDarkCurve(bri);
	const int noise = prm.detect_color_noise;
	auto set = [&](int xi, int yi) { flags[(size_t)yi * w + xi] = 1; };
	if(aaDifference(c, aaNormalized(film, w, x + 1, y), noise) >= thr) { set(x, y); set(x + 1, y); }
	if(aaDifference(c, aaNormalized(film, w, x, y + 1), noise) >= thr) { set(x, y); set(x, y + 1); }
	if(aaDifference(c, aaNormalized(film, w, x + 1, y + 1), noise) >= thr) { set(x, y); set(x + 1, y + 1); }
	if(x > 0 && aaDifference(c, aaNormalized(film, w, x - 1, y + 1), noise) >= thr) { set(x, y); set(x - 1, y + 1); }
	if(prm.variance_pixels > 0)
	{
		const int half = prm.variance_edge_size / 2;
		int vx = 0, vy = 0;
		for(int xd = -half; xd < half - 1; ++xd)
		{
			int xi = x + xd;
			if(xi < 0) xi = 0; else if(xi >= w - 1) xi = w - 2;
			if(aaDifference(aaNormalized(film, w, xi, y), aaNormalized(film, w, xi + 1, y), noise) >= thr) ++vx;
		}
		for(int yd = -half; yd < half - 1; ++yd)
		{
			int yi = y + yd;
			if(yi < 0) yi = 0; else if(yi >= h - 1) yi = h - 2;
			if(aaDifference(aaNormalized(film, w, x, yi), aaNormalized(film, w, x, yi + 1), noise) >= thr) ++vy;
		}
		if(vx + vy >= prm.variance_pixels)
			for(int xd = -half; xd < half; ++xd)
				for(int yd = -half; yd < half; ++yd)
				{
					int xi = x + xd, yi = y + yd;
					if(xi < 0) xi = 0; else if(xi >= w) xi = w - 1;
					if(yi < 0) yi = 0; else if(yi >= h) yi = h - 1;
					set(xi, yi);
				}
	}
}

// the decision over `film` (w x h x 5 floats on c's device), mask and count to the host; flags of different pixels only ever store 1, so the order is free
int nextPassMask(yafgpu_ctx *c, const float *film, const yafgpu_aa_noise_params *prm, std::vector<uint8_t> &mask, int64_t *n_resample)
{
	const int w = c->film.w, h = c->film.h;
	if(!(prm->threshold > 0.f)) { mask.clear(); if(n_resample) *n_resample = (int64_t)w * h; return YAFGPU_OK; }
	if(prm->dark_detection_type < YAFGPU_DARK_NONE || prm->dark_detection_type > YAFGPU_DARK_CURVE || prm->variance_edge_size < 0 || prm->variance_edge_size > 64)
		return setErr(c, YAFGPU_ERR_INVALID, "yafgpu_next_pass: dark_detection_type / variance_edge_size (0..64)");
	unsigned char *d_flags = nullptr;
	const size_t n = (size_t)w * h;
	YG_CUDA(c, poolAlloc(&d_flags, n, c->stream));
	cudaError_t e = cudaMemsetAsync(d_flags, 0, n, c->stream);
	if(e == cudaSuccess && w > 1 && h > 1)
	{
		k_next_pass<<<dim3((unsigned)((w + 255) / 256), (unsigned)h), 256, 0, c->stream>>>(film, w, h, *prm, d_flags);
		e = cudaGetLastError();
	}
	mask.resize(n);
	if(e == cudaSuccess) e = cudaMemcpyAsync(mask.data(), d_flags, n, cudaMemcpyDeviceToHost, c->stream);
	if(e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
	poolFree(d_flags, c->stream);
	if(e != cudaSuccess) return cudaErr(c, e, "yafgpu_next_pass");
	int64_t cnt = 0;
	for(size_t i = 0; i < n; ++i) cnt += mask[i] != 0;
	if(n_resample) *n_resample = cnt;
	return YAFGPU_OK;
}
} // namespace yg

extern "C" {

int yafgpu_next_pass(yafgpu_ctx *c, const yafgpu_aa_noise_params *prm, int64_t *n_resample)
{
	if(!c || !prm) return setErr(c, YAFGPU_ERR_INVALID, "NULL argument");
	if(!c->uploaded) return setErr(c, YAFGPU_ERR_STATE, "no scene uploaded");
	YG_CUDA(c, cudaSetDevice(c->device));
	c->aa_mask_valid = false;
	const int rc = nextPassMask(c, c->film.pix, prm, c->aa_mask, n_resample);
	if(rc == YAFGPU_OK) c->aa_mask_valid = true;
	return rc;
}

int yafgpu_host_register(void *ptr, size_t bytes)
{
	if(!ptr || !bytes) return setErr(nullptr, YAFGPU_ERR_INVALID, "yafgpu_host_register: NULL / empty buffer");
	const cudaError_t e = cudaHostRegister(ptr, bytes, cudaHostRegisterPortable);
	if(e == cudaErrorHostMemoryAlreadyRegistered) { cudaGetLastError(); return YAFGPU_OK; }
	return e == cudaSuccess ? YAFGPU_OK : cudaErr(nullptr, e, "cudaHostRegister");
}
int yafgpu_host_unregister(void *ptr)
{
	if(!ptr) return setErr(nullptr, YAFGPU_ERR_INVALID, "yafgpu_host_unregister: NULL");
	const cudaError_t e = cudaHostUnregister(ptr);
	if(e != cudaSuccess) cudaGetLastError();   // not sticky: the next launch must not inherit it
	return e == cudaSuccess ? YAFGPU_OK : cudaErr(nullptr, e, "cudaHostUnregister");
}

int yafgpu_sync(yafgpu_ctx *c)
{
	if(!c) return YAFGPU_ERR_INVALID;
	YG_CUDA(c, cudaSetDevice(c->device));
	return faultCheck(c, true);
}

int yafgpu_read_film(yafgpu_ctx *c, float *out)
{
	if(!c || !out) return setErr(c, YAFGPU_ERR_INVALID, "NULL argument");
	if(!c->uploaded) return setErr(c, YAFGPU_ERR_STATE, "no scene uploaded");
	YG_CUDA(c, cudaSetDevice(c->device));
	YG_CUDA(c, cudaMemcpyAsync(out, c->film.pix, sizeof(float) * 5 * (size_t)c->film.w * c->film.h, cudaMemcpyDeviceToHost, c->stream));
	YG_CUDA(c, cudaStreamSynchronize(c->stream));
	return YAFGPU_OK;
}

int yafgpu_write_film(yafgpu_ctx *c, const float *in)
{
	if(!c || !in) return setErr(c, YAFGPU_ERR_INVALID, "NULL argument");
	if(!c->uploaded) return setErr(c, YAFGPU_ERR_STATE, "no scene uploaded");
	YG_CUDA(c, cudaSetDevice(c->device));
	YG_CUDA(c, cudaMemcpyAsync(c->film.pix, in, sizeof(float) * 5 * (size_t)c->film.w * c->film.h, cudaMemcpyHostToDevice, c->stream));
	YG_CUDA(c, cudaStreamSynchronize(c->stream));
	return YAFGPU_OK;
}

int yafgpu_film_device_ptr(yafgpu_ctx *c, void **p, size_t *n)
{
	if(!c || !p || !n) return setErr(c, YAFGPU_ERR_INVALID, "NULL argument");
	if(!c->uploaded) return setErr(c, YAFGPU_ERR_STATE, "no scene uploaded");
	*p = c->film.pix; *n = (size_t)5 * c->film.w * c->film.h;
	return YAFGPU_OK;
}

int yafgpu_set_aa_multipliers(yafgpu_ctx *c, float light_mult, float indirect_mult)
{
	if(!c || !c->uploaded) return setErr(c, YAFGPU_ERR_STATE, "no scene uploaded");
	c->scene.integ.aa_light_sample_multiplier = light_mult; c->scene.integ.aa_indirect_sample_multiplier = indirect_mult;
	return YAFGPU_OK;
}

int yafgpu_set_option(yafgpu_ctx *c, const char *name, long long value)
{
	if(!c || !name) return setErr(c, YAFGPU_ERR_INVALID, "NULL argument");
	const std::string n(name);
	if(n == "wave_size") { if(value < 1024 || value > (1 << 24)) return setErr(c, YAFGPU_ERR_INVALID, "wave_size out of range"); c->wave_cap_override = (int)value; return YAFGPU_OK; }
	if(n == "leaf_size") { if(value < 1 || value > 8) return setErr(c, YAFGPU_ERR_INVALID, "leaf_size must be 1..8"); c->leaf_max = (int)value; return YAFGPU_OK; }
	if(n == "shade_streams") { if(value < 1 || value > 3) return setErr(c, YAFGPU_ERR_INVALID, "shade_streams must be 1..3"); c->shade_streams = (int)value; return YAFGPU_OK; }
	if(n == "ray_regeneration") { if(value < -1 || value > 3) return setErr(c, YAFGPU_ERR_INVALID, "ray_regeneration: -1..3"); c->ray_regeneration = (int)value; return YAFGPU_OK; }
	if(n == "regen_threshold") { if(value < 1 || value > 32) return setErr(c, YAFGPU_ERR_INVALID, "regen_threshold: 1..32"); c->regen_threshold = (int)value; return YAFGPU_OK; }
	if(n == "debug_stack_limit") { if(value < 0 || value > YG_STACK) return setErr(c, YAFGPU_ERR_INVALID, "debug_stack_limit: 0..96"); c->stack_limit = (int)value; return YAFGPU_OK; }
	if(n == "adaptive_iterations") { c->adaptive_iters = value != 0; return YAFGPU_OK; }
	if(n == "keep_nonfinite_samples") { c->keep_nonfinite = value != 0; return YAFGPU_OK; }
	if(n == "queue_fan") { if(value < 1 || value > 64) return setErr(c, YAFGPU_ERR_INVALID, "queue_fan out of range"); c->task_fan_override = (int)value; return YAFGPU_OK; }
	return setErr(c, YAFGPU_ERR_INVALID, "unknown option " + n);
}

int yafgpu_get_stats(yafgpu_ctx *c, yafgpu_stats *out)
{
	if(!c || !out) return setErr(c, YAFGPU_ERR_INVALID, "NULL argument");
	YG_CUDA(c, cudaSetDevice(c->device));
	YG_CUDA(c, cudaStreamSynchronize(c->stream));
	unsigned long long h[CNT_N];
	YG_CUDA(c, cudaMemcpy(h, c->d_counters, sizeof(h), cudaMemcpyDeviceToHost));
	*out = c->stats;
	out->rays_camera += h[CNT_RAYS_CAMERA]; out->rays_extension += h[CNT_RAYS_EXTENSION]; out->rays_shadow += h[CNT_RAYS_SHADOW];
	out->rays_shadow_transp += h[CNT_RAYS_SHADOW_TRANSP]; out->rays_light_probe += h[CNT_RAYS_LIGHT_PROBE];
	out->camera_samples += h[CNT_CAMERA_SAMPLES]; out->nodes_visited = h[CNT_NODES]; out->tris_tested = h[CNT_TRIS];
	out->nodes_visited_shadow = h[CNT_NODES_SHADOW]; out->tris_tested_shadow = h[CNT_TRIS_SHADOW]; out->nonfinite_samples = h[CNT_NONFINITE];
	return YAFGPU_OK;
}

int yafgpu_reset_stats(yafgpu_ctx *c)
{
	if(!c) return YAFGPU_ERR_INVALID;
	YG_CUDA(c, cudaSetDevice(c->device));
	YG_CUDA(c, cudaStreamSynchronize(c->stream));
	const yafgpu_stats old = c->stats;
	memset(&c->stats, 0, sizeof(c->stats));
	c->stats.bvh_nodes = old.bvh_nodes; c->stats.bvh_bytes = old.bvh_bytes; c->stats.tri_bytes = old.tri_bytes; c->stats.ms_build = old.ms_build;
	c->stats.bvh_node_bytes = old.bvh_node_bytes; c->stats.bvh_width = old.bvh_width; c->stats.bvh_levels = old.bvh_levels;
	YG_CUDA(c, cudaMemset(c->d_counters, 0, sizeof(unsigned long long) * CNT_N));
	return YAFGPU_OK;
}

} // extern "C"
