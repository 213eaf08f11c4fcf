// Wavefront integrator: K2 raygen, K3 closest-hit trace, K5-K8 shading (surface setup, light sampling, BSDFs, path logic),
// K4 shadow trace and K9 film accumulation.  Replaces, for one render pass,
//   TiledIntegrator::renderPass/renderTile      /root/reference/src/integrator/integrator_tiled.cc:261-521
//   PathIntegrator::integrate                   /root/reference/src/integrator/integrator_path_tracer.cc:112-347
//   DirectLightIntegrator::integrate            /root/reference/src/integrator/integrator_direct_light.cc:97-194
//   MonteCarloIntegrator::estimate*DirectLight,
//     doLightEstimation, recursiveRaytrace      /root/reference/src/integrator/integrator_montecarlo.cc:47-345,782-1028
//   Scene::isShadowed (opaque / transparent)    /root/reference/src/common/scene.cc:962-1035
//   ImageFilm::addSample                        /root/reference/src/common/imagefilm.cc:925-1015
//
// The reference evaluates one camera sample at a time as a recursion TREE (integrate -> recursiveRaytrace -> integrate ...,
// plus n_paths independent paths per integrate call).  Here the tree is flattened into a work list: every node is a task
// carrying the RGB factor its result is multiplied with on the way up to the camera sample (`slot`), so siblings and
// children can be shaded in any order and in parallel; results are summed into acc[slot] with float atomics and each
// camera sample is splatted into the film once, at the end of its wave.
//
// Deliberate deviations from the reference — all concern state that is schedule dependent on the CPU:
//   * Russian roulette draws come from a hash of (camera sample, path, depth) instead of the per-tile rand()-seeded MWC;
//   * estimateOneDirectLight picks its light from a hash instead of the per-thread Halton counter;
//   * RenderState::include_lights_ is "true" for every integrate() task (the reference leaves it at whatever the last
//     path of a sibling subtree wrote, integrator_path_tracer.cc:212,253).
#include "context.h"
#include "lights.cuh"
#include "nodes.cuh"
#include <algorithm>
#include <cstdio>

namespace yg {


struct WaveParams
{
	int n_slots;        // camera samples in this wave = n_pix * n_smp
	int pix_base, n_pix; // pixel-list range of the wave
	int sample_base;    // first sample index of the wave inside the pass
	int n_samples_pass; // samples per pixel of the whole pass (integrator_tiled.cc:385,400)
	int pass_offs;      // sampling offset of the pass (integrator_tiled.cc:335)
	int aa_passes;
	int task_cap, shadow_cap;
	unsigned seed;
};

// ---- packed task state (wavefront.h) ---------------------------------------------------------------------------------
__device__ __forceinline__ unsigned packState(int kind, int raylevel, int depth, int ray_division, int caustic, int add_depth)
{
	return (unsigned)kind | ((unsigned)raylevel << 1) | ((unsigned)depth << 6) | ((unsigned)min(ray_division, 255) << 12) | ((unsigned)caustic << 22) |
	       ((unsigned)min(add_depth, 31) << 23);
}
__device__ __forceinline__ int stKind(unsigned p) { return p & 1; }
__device__ __forceinline__ int stRaylevel(unsigned p) { return (p >> 1) & 31; }
__device__ __forceinline__ int stDepth(unsigned p) { return (p >> 6) & 63; }
__device__ __forceinline__ int stDivision(unsigned p) { return (p >> 12) & 255; }
__device__ __forceinline__ int stCaustic(unsigned p) { return (p >> 22) & 1; }
__device__ __forceinline__ int stAddDepth(unsigned p) { return (p >> 23) & 31; }
// bit 20: RenderState::chromatic_ == false — the task carries ONE wavelength (st.y of a TASK_INTEGRATE holds its bits), set by the dispersive
// branch of recursiveRaytrace and inherited by everything below it
#define YG_ST_MONO (1u << 20)
// bit 21: this integrate() was called by the dispersive branch, which sets RenderState::include_lights_ = false for its children
// (integrator_montecarlo.cc:799 "debatable..."); the glossy / specular branches set it back to true for theirs (:864,977)
#define YG_ST_NOLIGHTS (1u << 21)
// bit 28 ("bg_transp_refract"): this integrate() decides the alpha of its camera sample.  integrate() returns alpha = m + (1 - m) * alpha_child with
// m = Material::getAlpha and alpha_child = what the LAST refracted child returned (integrator_path_tracer.cc:319-323, integrator_montecarlo.cc:954,1023);
// unrolled from the camera this is alpha = A + B * alpha_here, and the task carries (A, B) in thr.xy, which a TASK_INTEGRATE does not use otherwise.
#define YG_ST_ALPHA (1u << 28)
// acc[slot].w holds the alpha of the camera sample as float bits XOR 1.0f's, so that the zero the wave starts from reads as alpha = 1
__device__ __forceinline__ float alphaEncode(float alpha) { return __int_as_float(__float_as_int(alpha) ^ 0x3f800000); }

// warp-aggregated queue slot allocation
__device__ __forceinline__ int allocSlot(int *counter)
{
	const unsigned m = __activemask();
	const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
	int base = 0;
	if(lane == leader) base = atomicAdd(counter, __popc(m));
	base = __shfl_sync(m, base, leader);
	return base + __popc(m & ((1u << lane) - 1u));
}

__device__ __forceinline__ unsigned hash3(unsigned a, unsigned b, unsigned c)
{
	unsigned h = a * 0x9E3779B1u;
	h ^= b + 0x7F4A7C15u + (h << 6) + (h >> 2);
	h *= 0x85EBCA6Bu; h ^= h >> 13;
	h ^= c + 0x165667B1u + (h << 6) + (h >> 2);
	h *= 0xC2B2AE35u; h ^= h >> 16;
	h *= 0x27D4EB2Fu; h ^= h >> 15;
	return h;
}
__device__ __forceinline__ float hashRand(unsigned a, unsigned b, unsigned c) { return (float)((double)hash3(a, b, c) * YG_MULT_RATIO); } // Random::operator() is x * 2^-32

#ifdef YG_NAN_TRAP   // diagnosis build (tools/nan_probe.py): name the source line of a non-finite contribution
#define accAdd(a, s, c) accAddAt(a, s, c, __LINE__)
__device__ __forceinline__ void accAddAt(float4 *acc, int slot, Col c, int line)
{
	if(!(isfinite(c.r) && isfinite(c.g) && isfinite(c.b))) printf("non-finite contribution at integrator.cu:%d slot %d (%g %g %g)\n", line, slot, c.r, c.g, c.b);
#else
__device__ __forceinline__ void accAdd(float4 *acc, int slot, Col c)
{
#endif
	float *p = (float *)(acc + slot);
	if(c.r != 0.f) atomicAdd(p, c.r);
	if(c.g != 0.f) atomicAdd(p + 1, c.g);
	if(c.b != 0.f) atomicAdd(p + 2, c.b);
}

struct SlotInfo { int x, y, sample; unsigned sampling_offs; int pixel_sample; };
__device__ __forceinline__ SlotInfo slotInfo(const WaveParams &wp, const uint32_t *__restrict__ pixels, int slot)
{
	SlotInfo s;
	const int pi = slot % wp.n_pix, si = slot / wp.n_pix;
	const uint32_t pxy = __ldg(pixels + wp.pix_base + pi);
	s.x = (int)(pxy & 0xffffu); s.y = (int)(pxy >> 16);
	s.sample = wp.sample_base + si;
	s.sampling_offs = fnv32a((uint32_t)s.y * fnv32a((uint32_t)s.x)); // fnv32ABuf__(i * fnv32ABuf__(j)) integrator_tiled.cc:379
	s.pixel_sample = wp.pass_offs + s.sample;                          // :389
	return s;
}
// pixel-sample position integrator_tiled.cc:392-403
__device__ __forceinline__ void samplePos(const WaveParams &wp, const SlotInfo &si, float &dx, float &dy)
{
	dx = 0.5f; dy = 0.5f;
	if(wp.aa_passes > 1)
	{
		dx = riVdC((uint32_t)si.pixel_sample, si.sampling_offs);
		dy = riS((uint32_t)si.pixel_sample, si.sampling_offs);
	}
	else if(wp.n_samples_pass > 1)
	{
		const float d_1 = (float)(1.0 / (double)(float)wp.n_samples_pass);
		dx = (float)((0.5 + (double)(float)si.sample) * (double)d_1);
		dy = riLp((uint32_t)si.sample + si.sampling_offs);
	}
}

// ---- K2: camera rays ---------------------------------------------------------------------------------------------------
__device__ void shirleyDisk(float r_1, float r_2, float &u, float &v);

// PerspectiveCamera::biasDist / sampleTsd / getLensUv camera_perspective.cc:76-128
__device__ __forceinline__ float biasDist(int bias, float r)
{
	if(bias == YAFGPU_BB_CENTER) return fSqrt(fSqrt(r) * r);
	if(bias == YAFGPU_BB_EDGE) return fSqrt((float)1.0 - r * r);
	return fSqrt(r);
}
__device__ __forceinline__ void lensUv(const CameraDev &c, float r_1, float r_2, float &u, float &v)
{
	switch(c.bokeh_type)
	{
		case YAFGPU_BK_TRI: case YAFGPU_BK_SQR: case YAFGPU_BK_PENTA: case YAFGPU_BK_HEXA:
		{
			const float fn = (float)c.bokeh_type;
			int idx = (int)(r_1 * fn);
			r_1 = (r_1 - ((float)idx) / fn) * fn;
			r_1 = biasDist(c.bokeh_bias, r_1);
			const float b_1 = r_1 * r_2, b_0 = r_1 - b_1;
			idx <<= 1;
			u = c.ls[idx] * b_0 + c.ls[idx + 2] * b_1;
			v = c.ls[idx + 1] * b_0 + c.ls[idx + 3] * b_1;
			break;
		}
		case YAFGPU_BK_DISK2: case YAFGPU_BK_RING:
		{
			const float w = (float)YG_M_2PI * r_2;
			if(c.bokeh_type == YAFGPU_BK_RING) r_1 = fSqrt((float)0.707106781 + (float)0.292893218);
			else r_1 = biasDist(c.bokeh_bias, r_1);
			u = r_1 * fCos(w);
			v = r_1 * fSin(w);
			break;
		}
		default: shirleyDisk(r_1, r_2, u, v);
	}
}

// shirleyDisk__ util_sample.h (concentric map) used by the default bokeh
__device__ void shirleyDisk(float r_1, float r_2, float &u, float &v)
{
	float phi = 0.f, r = 0.f;
	const float a = 2.f * r_1 - 1.f, b = 2.f * r_2 - 1.f;
	if(a > -b)
	{
		if(a > b) { r = a; phi = (float)(0.78539816339744830962 * (double)(b / a)); }
		else { r = b; phi = (float)(0.78539816339744830962 * (double)(2.f - a / b)); }
	}
	else
	{
		if(a < b) { r = -a; phi = (float)(0.78539816339744830962 * (double)(4.f + b / a)); }
		else { r = -b; if(b != 0) phi = (float)(0.78539816339744830962 * (double)(6.f - a / b)); else phi = 0.f; }
	}
	u = r * fCos(phi); v = r * fSin(phi);
}

__global__ void __launch_bounds__(256) k_raygen(SceneDev sc, WaveParams wp, const uint32_t *__restrict__ pixels, TaskQueue q)
{
	for(int slot = blockIdx.x * blockDim.x + threadIdx.x; slot < wp.n_slots; slot += gridDim.x * blockDim.x)
	{
		const SlotInfo si = slotInfo(wp, pixels, slot);
		float dx, dy;
		samplePos(wp, si, dx, dy);
		const CameraDev &c = sc.cam;
		const float px = (float)si.x + dx, py = (float)si.y + dy;
		V3 from = ldv(c.pos), dir;
		bool dead = false; // Camera::shootRay returned wt = 0: the film gets a black sample (integrator_tiled.cc:412-416)
		switch(c.type)
		{
			case YAFGPU_CAM_ORTHOGRAPHIC: // OrthographicCamera::shootRay camera_orthographic.cc:53-64
				from = (from + ldv(c.vright) * px) + ldv(c.vup) * py;
				dir = ldv(c.cam_to);
				break;
			case YAFGPU_CAM_ANGULAR: // AngularCamera::shootRay camera_angular.cc:57-85 (asin/atan/atan2 are the double overloads there)
			{
				const float u = 1.f - 2.f * (px / (float)c.resx);
				float v = 2.f * (py / (float)c.resy) - 1.f;
				v *= c.aspect_ratio;
				const float radius = fSqrt(u * u + v * v);
				dir = ldv(c.cam_to);
				if(c.circular && radius > c.max_radius) { dead = true; break; }
				float theta = 0.f;
				if(!((u == 0) && (v == 0))) theta = (float)atan2((double)v, (double)u);
				float phi;
				if(c.projection == YAFGPU_AP_ORTHOGRAPHIC) phi = (float)asin((double)(radius / c.focal_length));
				else if(c.projection == YAFGPU_AP_STEREOGRAPHIC) phi = (float)((double)2.f * atan((double)(radius / (2.f * c.focal_length))));
				else if(c.projection == YAFGPU_AP_EQUISOLID) phi = (float)((double)2.f * asin((double)(radius / (2.f * c.focal_length))));
				else if(c.projection == YAFGPU_AP_RECTILINEAR) phi = (float)atan((double)(radius / c.focal_length));
				else phi = radius / c.focal_length;
				dir = fSin(phi) * (fCos(theta) * ldv(c.vright) + fSin(theta) * ldv(c.vup)) + fCos(phi) * ldv(c.cam_to);
				break;
			}
			case YAFGPU_CAM_EQUIRECTANGULAR: // EquirectangularCamera::shootRay camera_equirectangular.cc:50-66
			{
				const float u = 2.f * px / (float)c.resx - 1.f;
				const float v = 2.f * py / (float)c.resy - 1.f;
				const float phi = (float)(YG_M_PI * (double)u);
				const float theta = (float)(1.57079632679489661923 * (double)v);
				dir = fCos(theta) * (fCos(phi) * ldv(c.cam_to) + fSin(phi) * ldv(c.vright)) + fSin(theta) * ldv(c.vup);
				break;
			}
			default: // PerspectiveCamera::shootRay camera_perspective.cc:133-155 (ArchitectCamera inherits it)
				dir = normalize((ldv(c.vright) * px + ldv(c.vup) * py) + ldv(c.vto));
		}
		const V3 nn = ldv(c.near_n);
		// rayPlaneIntersection__ util_geometry.h:34-37 (near and far plane share the normal)
		const float tmin = dot(nn, ldv(c.near_p) - from) / dot(dir, nn);
		float tmax = dot(nn, ldv(c.far_p) - from) / dot(dir, nn);
		if(c.aperture != 0)
		{
			HaltonB<3> hu; HaltonB<5> hv;
			hu.setStart((uint32_t)wp.pass_offs + si.sampling_offs + (uint32_t)si.sample);
			hv.setStart((uint32_t)wp.pass_offs + si.sampling_offs + (uint32_t)si.sample);
			const float lu = hu.getNext(), lv = hv.getNext();
			float u, v;
			lensUv(c, lu, lv, u, v);
			const V3 li = ldv(c.dof_rt) * u + ldv(c.dof_up) * v;
			from = from + li;
			dir = normalize((dir * c.dof_distance) - li);
		}
		// a dead sample travels as a zero-length ray of zero weight: it misses everything and adds nothing, k_film_add still counts its filter weight
		const float w0 = dead ? 0.f : 1.f;
		if(dead) tmax = 0.f;
		q.ray_o[slot] = make_float4(from.x, from.y, from.z, tmin);
		q.ray_d[slot] = make_float4(dir.x, dir.y, dir.z, tmax);
		q.wgt[slot] = make_float4(w0, w0, w0, 0.f);
		const bool alpha_chain = sc.integ.transp_refr_bg != 0;
		q.thr[slot] = alpha_chain ? make_float4(0.f, 1.f, 1.f, 0.f) : make_float4(1.f, 1.f, 1.f, 0.f);
		q.st[slot] = make_uint4((unsigned)slot, 0u, packState(TASK_INTEGRATE, 0, 0, 1, 0, 0) | (alpha_chain ? YG_ST_ALPHA : 0u), 0u);
		q.dc[slot] = make_float2(0.f, 0.f);
	}
}

// ---- K3: closest hit for every queued task -----------------------------------------------------------------------------
// Besides the hit record the kernel bins every task by (task kind, material) — the key of the counting sort that makes
// warps of the shading kernel run ONE material's code (the reference dispatches per hit through a virtual call; here
// divergence inside a warp is the cost that matters, ncu: 2.6 active threads per instruction without the sort).
struct SortBins
{
	const int *tri_mat;     // per global triangle: material index
	const int *mat_bucket;  // per material: rank in (type, id) order
	int n_mat;              // buckets: kind * n_mat + mat_bucket, last bucket (2 * n_mat) = miss
	int *key;               // per task: bucket
	int *hist;              // per bucket: count (zeroed before the launch)
};

// warp-aggregated atomicAdd(&counter[key], 1) for the lanes named in `valid`; returns this lane's slot
__device__ __forceinline__ int binAdd(int *counter, int key, unsigned valid)
{
	const unsigned m = __match_any_sync(valid, key);
	const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
	int base = 0;
	if(lane == leader) base = atomicAdd(counter + key, __popc(m));
	base = __shfl_sync(m, base, leader);
	return base + __popc(m & ((1u << lane) - 1u));
}

struct TraceIO
{
	TaskQueue q; SortBins bins;
	__device__ __forceinline__ void load(int i, V3 &o, V3 &d, float &tmin, float &dist) const
	{
		const float4 a = q.ray_o[i], b = q.ray_d[i];
		o = mk(a.x, a.y, a.z); tmin = a.w; d = mk(b.x, b.y, b.z);
		dist = (b.w < 0.f) ? __int_as_float(0x7f800000) : b.w; // scene.cc:900-901
	}
	__device__ __forceinline__ void store(int i, const HitD &h, unsigned lanes) const
	{
		q.hit[i] = make_float4(h.t, h.u, h.v, __int_as_float(h.tri));
		const int kind = (int)(q.st[i].z & 1u);
		const int key = h.tri < 0 ? 2 * bins.n_mat : kind * bins.n_mat + __ldg(bins.mat_bucket + __ldg(bins.tri_mat + h.tri));
		bins.key[i] = key;
		binAdd(bins.hist, key, lanes);
	}
};

// task i -> thread i (grid-stride): the warp's 32 consecutive tasks stay together — queue order is spatially coherent (tile order, then
// per-material runs), which is worth more than full lanes once the BVH exceeds the caches (profiles/r1_traversal.md).
// REGEN: warp-persistent schedule with ray regeneration (traverse.cuh traversePersistent) instead of task i -> thread i.
template<bool COUNT, bool REGEN>
__global__ void __launch_bounds__(128) k_trace(Accel acc, TaskQueue q, const int *__restrict__ n_tasks, int task_cap, SortBins bins, int *__restrict__ next, int refill_below,
                                               unsigned long long *__restrict__ counters, const WaveCtl *__restrict__ ctl)
{
	if(ctl->overflow) return;   // a queue ran over earlier in this wave: the wave is discarded, do not pay for the rest of it
	const int n = min(*n_tasks, task_cap);
	TravCounters tc; tc.nodes = 0; tc.tris = 0;
	TraceIO io; io.q = q; io.bins = bins;
	__shared__ uint2 s_stack[YG_STACK_SHARED * 128];
	TravStack<YG_STACK_SHARED> st; bindSharedStack(st, s_stack);
	// bucket counts go to a per-block shared histogram first: with a handful of materials every warp of the grid would otherwise
	// hit the same few L2 addresses (2.6e5 same-address atomics per 8 M tasks serialise in one L2 slice)
	extern __shared__ int s_hist[];
	const int nb = 2 * bins.n_mat + 1;
	const bool block_hist = nb <= YG_SHARED_BUCKETS;
	if(block_hist)
	{
		for(int b = threadIdx.x; b < nb; b += blockDim.x) s_hist[b] = 0;
		__syncthreads();
		io.bins.hist = s_hist;
	}
	if(REGEN) traversePersistent<false, COUNT>(acc, n, next, io, st, &tc, refill_below);
	else
	{
		const int stride = gridDim.x * blockDim.x;
		for(int base = blockIdx.x * blockDim.x + (threadIdx.x & ~31); base < n; base += stride)
		{
			const int i = base + (threadIdx.x & 31);
			const bool valid = i < n;
			const unsigned vmask = __ballot_sync(0xffffffffu, valid);
			if(!valid) continue;
			RayD r;
			float dist;
			io.load(i, r.o, r.d, r.tmin, dist);
			r.tmax = dist;
			io.store(i, traceClosestS<COUNT>(acc, r, st, &tc), vmask);
		}
	}
	if(block_hist)
	{
		__syncthreads();
		for(int b = threadIdx.x; b < nb; b += blockDim.x) if(s_hist[b]) atomicAdd(bins.hist + b, s_hist[b]);
	}
	if(COUNT) { atomicAdd(&counters[CNT_NODES], tc.nodes); atomicAdd(&counters[CNT_TRIS], tc.tris); }
}

// exclusive scan of the bucket counts (one block; at most a few thousand buckets)
__global__ void __launch_bounds__(1024) k_bin_scan(const int *__restrict__ hist, int *__restrict__ offs, int nb)
{
	__shared__ int part[1024];
	const int per = (nb + 1023) / 1024;
	const int b0 = threadIdx.x * per;
	int sum = 0;
	for(int k = 0; k < per; ++k) if(b0 + k < nb) sum += hist[b0 + k];
	part[threadIdx.x] = sum;
	__syncthreads();
	for(int d = 1; d < 1024; d <<= 1)
	{
		const int v = threadIdx.x >= d ? part[threadIdx.x - d] : 0;
		__syncthreads();
		part[threadIdx.x] += v;
		__syncthreads();
	}
	int run = part[threadIdx.x] - sum;
	for(int k = 0; k < per; ++k) if(b0 + k < nb) { offs[b0 + k] = run; run += hist[b0 + k]; }
}

// scatter: perm[slot in bucket order] = task index.  A block ranks a chunk of YG_SCATTER_PER_THREAD * 256 tasks in shared memory and
// reserves one range per non-empty bucket, so the global offsets see a few atomics per 4096 tasks instead of one or more per warp.
#define YG_SCATTER_PER_THREAD 16
__global__ void __launch_bounds__(256) k_bin_scatter_block(const int *__restrict__ n_tasks, int task_cap, const int *__restrict__ key, int *__restrict__ offs, int *__restrict__ perm, int nb)
{
	extern __shared__ int s_cnt[]; // [nb] counts, then [nb] reserved bases
	int *s_base = s_cnt + nb;
	const int n = min(*n_tasks, task_cap);
	const int chunk = YG_SCATTER_PER_THREAD * 256;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	for(int c0 = blockIdx.x * chunk; c0 < n; c0 += gridDim.x * chunk)
	{
		for(int b = threadIdx.x; b < nb; b += 256) s_cnt[b] = 0;
		__syncthreads();
		int k[YG_SCATTER_PER_THREAD], r[YG_SCATTER_PER_THREAD];
#pragma unroll
		for(int j = 0; j < YG_SCATTER_PER_THREAD; ++j)
		{	// warp w takes 32 consecutive tasks per round: ranks inside a bucket follow queue order up to the warp interleave
			const int i = c0 + (j * 8 + warp) * 32 + lane;
			const bool valid = i < n;
			const unsigned vmask = __ballot_sync(0xffffffffu, valid);
			k[j] = -1; r[j] = 0;
			if(valid) { k[j] = key[i]; r[j] = binAdd(s_cnt, k[j], vmask); }
		}
		__syncthreads();
		for(int b = threadIdx.x; b < nb; b += 256) if(s_cnt[b]) s_base[b] = atomicAdd(offs + b, s_cnt[b]);
		__syncthreads();
#pragma unroll
		for(int j = 0; j < YG_SCATTER_PER_THREAD; ++j)
			if(k[j] >= 0) perm[s_base[k[j]] + r[j]] = c0 + (j * 8 + warp) * 32 + lane;
		__syncthreads();
	}
}
__global__ void __launch_bounds__(256) k_bin_scatter(const int *__restrict__ n_tasks, int task_cap, const int *__restrict__ key, int *__restrict__ offs, int *__restrict__ perm)
{
	const int n = min(*n_tasks, task_cap);
	const int stride = gridDim.x * blockDim.x;
	for(int base = blockIdx.x * blockDim.x + (threadIdx.x & ~31); base < n; base += stride)
	{
		const int i = base + (threadIdx.x & 31);
		const bool valid = i < n;
		const unsigned vmask = __ballot_sync(0xffffffffu, valid);
		if(!valid) continue;
		perm[binAdd(offs, key[i], vmask)] = i;
	}
}

// ---- shading helpers ---------------------------------------------------------------------------------------------------
struct ShadeCtx
{
	const SceneDev *sc;
	WaveParams wp;
	TaskQueue out;
	ShadowQueue sq;
	float4 *acc;
	WaveCtl *ctl;
	int iter;
};

__device__ __forceinline__ void pushShadow(const ShadeCtx &cx, int slot, V3 from, float tmin, V3 dir, float tmax, Col contrib)
{
	if(isBlack(contrib)) return;
	const int k = allocSlot(&cx.ctl->n_shadow[cx.iter]);
	if(k >= cx.wp.shadow_cap) { cx.ctl->overflow = 1; return; }
	cx.sq.ray_o[k] = make_float4(from.x, from.y, from.z, tmin);
	cx.sq.ray_d[k] = make_float4(dir.x, dir.y, dir.z, tmax);
	cx.sq.contrib[k] = make_float4(contrib.r, contrib.g, contrib.b, __int_as_float(slot));
}

// `beer`: 1 + index of the absorbing material whose Beer transmittance over the traced segment multiplies everything the child task returns
// (recursiveRaytrace: `integ *= vcol` after integrate(), integrator_montecarlo.cc:841-845,912-915,991-994,1019-1022); 0 = none
#ifdef YG_NAN_TRAP
#define pushTask(cx, slot, from, tmin, dir, wgt, ...) pushTaskAt(__LINE__, cx, slot, from, tmin, dir, wgt, __VA_ARGS__)
__device__ __forceinline__ int pushTaskAt(int line, const ShadeCtx &cx, int slot, V3 from, float tmin, V3 dir, Col wgt, Col thr, unsigned offs, unsigned packed, unsigned ray_offset, float dc1, float dc2, int beer = 0)
{
	if(!(isfinite(wgt.r) && isfinite(wgt.g) && isfinite(wgt.b) && isfinite(thr.r) && isfinite(thr.g) && isfinite(thr.b) && isfinite(dir.x + dir.y + dir.z) && isfinite(from.x + from.y + from.z)))
		printf("non-finite task pushed at integrator.cu:%d slot %d wgt (%g %g %g) thr (%g %g %g) dir (%g %g %g) from (%g %g %g) packed %08x\n", line, slot, wgt.r, wgt.g, wgt.b, thr.r, thr.g, thr.b, dir.x, dir.y, dir.z, from.x, from.y, from.z, packed);
#else
__device__ __forceinline__ int pushTask(const ShadeCtx &cx, int slot, V3 from, float tmin, V3 dir, Col wgt, Col thr, unsigned offs, unsigned packed, unsigned ray_offset, float dc1, float dc2, int beer = 0)
{
#endif
	const int k = allocSlot(&cx.ctl->n_tasks[cx.iter + 1]);
	if(k >= cx.wp.task_cap) { cx.ctl->overflow = 1; return -1; }
	cx.out.ray_o[k] = make_float4(from.x, from.y, from.z, tmin);
	cx.out.ray_d[k] = make_float4(dir.x, dir.y, dir.z, -1.f);
	cx.out.wgt[k] = make_float4(wgt.r, wgt.g, wgt.b, __int_as_float(beer));
	cx.out.thr[k] = make_float4(thr.r, thr.g, thr.b, 0.f);
	cx.out.st[k] = make_uint4((unsigned)slot, offs, packed, ray_offset);
	cx.out.dc[k] = make_float2(dc1, dc2);
	return k;
}

// Material::getVolumeHandler(sp.ng_ * ref_ray.dir_ < 0): only the inside handler (vol_i_) exists, created by the glass factories for "absorption"
// (a compile-time `false` in the kernels of material types that cannot absorb)
#define YG_CAN_ABSORB(MT) ((MT) < 0 || (MT) == YAFGPU_MAT_GLASS || (MT) == YAFGPU_MAT_ROUGH_GLASS)
__device__ __forceinline__ int beerOfImpl(const MatDev &m, const SurfPt &sp, V3 dir) { return (m.absorb && dot(sp.ng, dir) < 0.f) ? sp.mat + 1 : 0; }
#define beerOf(m, sp, dir) (YG_CAN_ABSORB(MT) ? beerOfImpl(m, sp, dir) : 0)

// MonteCarloIntegrator::doLightEstimation integrator_montecarlo.cc:78-345 (Combined pass only).
// `W` is the factor the returned colour would be multiplied with before reaching the camera sample; every term is queued
// as a shadow ray carrying its finished contribution (or added at once when the light/material pair casts no shadow).
#ifndef YG_DLE_INLINE
#define YG_DLE_INLINE 1
#endif
#ifndef YG_SHADE_MINB
#define YG_SHADE_MINB 4
#endif
#ifndef YG_SHADE_MINB_L
#define YG_SHADE_MINB_L 4 /* resident blocks per SM of the diffuse-only kernel (fewer live values than the full BSDFs) */
#endif
#if YG_DLE_INLINE
#define YG_DLE YG_D
#else
#define YG_DLE YG_DN
#endif
template<int MT>
YG_DLE void doLightEstimation(const ShadeCtx &cx, const MatDev &m, int slot, const SlotInfo &si, int li, const SurfPt &sp, V3 wo, unsigned loffs, int ray_division, Col W)
{
	const SceneDev &sc = *cx.sc;
	const LightDev &L = sc.lights[li];
	const bool cast_shadows = L.cast_shadows && m.receive_shadows;
	const unsigned l_offs = loffs * 4567u;
	const float plen = length(sp.p);
	const float sh_tmin = sc.integ.shadow_bias_auto ? sc.integ.shadow_bias * fmaxf(1.f, plen) : sc.integ.shadow_bias;
	if(lightIsDirac(L))
	{
		Col lcol; V3 ldir; float tmax;
		if(lightIlluminate(sc, L, sp, lcol, ldir, tmax))
		{
			const float angle = m.flat ? 1.f : fabsf(dot(sp.n, ldir));
			const Col surf_col = matEval<MT>(sc, m, sp, wo, ldir, BsdfAll);
			const Col c = W * ((surf_col * lcol) * angle);
			if(cast_shadows) pushShadow(cx, slot, sp.p, sh_tmin, ldir, tmax, c);
			else accAdd(cx.acc, slot, c);
		}
		return;
	}
	int n = (int)ceilf((float)L.samples * sc.integ.aa_light_sample_multiplier);
	if(ray_division > 1) n = max(1, n / ray_division);
	const float inv_ns = 1.f / (float)n;
	const unsigned offs = (unsigned)n * (unsigned)si.pixel_sample + si.sampling_offs + l_offs;
	const bool can_intersect = lightCanIntersect(L) && !((L.type == YAFGPU_LIGHT_MESH || L.type == YAFGPU_LIGHT_BGPORTAL) && L.accel < 0);
	Halton2 hal_2; HaltonB<3> hal_3;
	hal_2.setStart(offs - 1u); hal_3.setStart(offs - 1u);
	const double hal_3_start = hal_3.value; // the BSDF loop below restarts both sequences at the same element
	for(int i = 0; i < n; ++i)
	{
		const float s1 = hal_2.getNext(), s2 = hal_3.getNext();
		Col lscol; float lpdf, tmax; V3 ldir;
		if(!lightIllumSample(sc, L, sp, s1, s2, lscol, lpdf, ldir, tmax)) continue;
		if(!(lpdf > 1e-6f)) continue;
		const Col surf_col = matEval<MT>(sc, m, sp, wo, ldir, BsdfAll);
		const float angle = m.flat ? 1.f : fabsf(dot(sp.n, ldir));
		Col c;
		if(can_intersect)
		{
			const float m_pdf = matPdf<MT>(sc, m, sp, wo, ldir, BsdfGlossy | BsdfDiffuse | BsdfDispersive | BsdfReflect | BsdfTransmit);
			if(m_pdf > 1e-6f)
			{
				const float l_2 = lpdf * lpdf, m_2 = m_pdf * m_pdf;
				const float w = l_2 / (l_2 + m_2);
				c = (((surf_col * lscol) * angle) * w) / lpdf;
			}
			else c = ((surf_col * lscol) * angle) / lpdf;
		}
		else c = ((surf_col * lscol) * angle) / lpdf;
		c = W * (c * inv_ns);
		if(cast_shadows) pushShadow(cx, slot, sp.p, sh_tmin, ldir, tmax, c);
		else accAdd(cx.acc, slot, c);
	}
	if(can_intersect)
	{
		hal_2.setStart(offs - 1u); hal_3.value = hal_3_start;
		const float b_tmin = sc.integ.ray_min_dist_auto ? sc.integ.ray_min_dist * fmaxf(1.f, plen) : sc.integ.ray_min_dist;
		for(int i = 0; i < n; ++i)
		{
			BSample s; s.s1 = hal_2.getNext(); s.s2 = hal_3.getNext(); s.pdf = 0.f; s.sampled = BsdfNone;
			s.flags = BsdfGlossy | BsdfDiffuse | BsdfDispersive | BsdfReflect | BsdfTransmit;
			float Wm = 0.f; V3 bdir = mk(0.f);
			const Col surf_col = matSample<MT>(sc, m, sp, wo, bdir, s, Wm);
			Col lcol; float light_pdf, t;
			if(s.pdf > 1e-6f && lightIntersect(sc, L, sp.p, bdir, b_tmin, t, lcol, light_pdf))
			{
				if(!(light_pdf > 1e-6f)) continue;
				const float l_pdf = 1.f / light_pdf;
				const float l_2 = l_pdf * l_pdf, m_2 = s.pdf * s.pdf;
				const float w = m_2 / (l_2 + m_2);
				const Col c = W * ((((surf_col * lcol) * w) * Wm) * inv_ns);
				if(cast_shadows) pushShadow(cx, slot, sp.p, b_tmin, bdir, t, c);
				else accAdd(cx.acc, slot, c);
			}
		}
	}
}

// ---- K5-K8: one thread per task, one kernel per material type -----------------------------------------------------------
// k_shade<MT> shades the tasks whose hit material is of type MT (MT_MISS: the rays that left the scene).  The counting sort
// leaves the tasks of one type in two contiguous runs of `perm` (one per task kind); `ends` holds the bucket end offsets the
// scatter kernel leaves behind, [b0, b1] is the bucket range of the type inside one kind (host-known), so a launch needs no
// host-side knowledge of the counts.  Compiling the material code in per type keeps each kernel's instruction footprint
// (and register allocation) to what that BSDF needs: the single generic kernel was 278 k SASS instructions.
enum { MT_MISS = 100, MT_NODES = -2 /* negative: the material type is read from the record, as for blend children */ };

// KIND: 0 = only the TASK_INTEGRATE run of the type's buckets, 1 = only the TASK_PATH run, 2 = both (each kernel then holds the code and
// the live state of one branch only)
template<int MT, int KIND>
__global__ void __launch_bounds__(128, MT == YG_MAT_SD_LAMBERT ? YG_SHADE_MINB_L : YG_SHADE_MINB) k_shade(const __grid_constant__ SceneDev sc, WaveParams wp, const uint32_t *__restrict__ pixels, TaskQueue in, TaskQueue out, ShadowQueue sq,
                                               float4 *__restrict__ acc, WaveCtl *__restrict__ ctl, int iter, const int *__restrict__ perm,
                                               const int *__restrict__ ends, int n_mat, int b0, int b1)
{
	if(ctl->overflow) return;
	ShadeCtx cx; cx.sc = &sc; cx.wp = wp; cx.out = out; cx.sq = sq; cx.acc = acc; cx.ctl = ctl; cx.iter = iter;
	const IntegratorDev &ig = sc.integ;
	// runs of this type: kind 0 buckets [b0, b1], kind 1 buckets [n_mat + b0, n_mat + b1]; the miss bucket is 2 * n_mat
	int start_a, len_a, start_b, len_b;
	if(MT == MT_MISS) { start_a = ends[2 * n_mat - 1]; len_a = ends[2 * n_mat] - start_a; start_b = 0; len_b = 0; }
	else
	{
		start_a = b0 > 0 ? ends[b0 - 1] : 0; len_a = ends[b1] - start_a;
		start_b = ends[n_mat + b0 - 1]; len_b = ends[n_mat + b1] - start_b;
		if(KIND == 0) len_b = 0;
		if(KIND == 1) len_a = 0;
	}
	const int n = len_a + len_b;
	for(int ii = blockIdx.x * blockDim.x + threadIdx.x; ii < n; ii += gridDim.x * blockDim.x)
	{
		const int i = perm[ii < len_a ? start_a + ii : start_b + (ii - len_a)];
		const float4 ro = in.ray_o[i], rd = in.ray_d[i], hh = in.hit[i], wg = in.wgt[i];
		const uint4 st = in.st[i];
		const int slot = (int)st.x;
		const unsigned packed = st.z;
		const V3 from = mk(ro.x, ro.y, ro.z), dir = mk(rd.x, rd.y, rd.z);
		Col W = mkc(wg.x, wg.y, wg.z);
		const int tri = __float_as_int(hh.w);
		if(const int beer = __float_as_int(wg.w))
		{	// BeerVolumeHandler::transmittance volumehandler_beer.cc:36-47 over the segment just traced; a ray that left the scene has
			// tmax < 0 there and the handler returns black, so nothing the child would add survives
			if(MT == MT_MISS || tri < 0 || hh.x > 1e30f) continue;
			const MatDev &bm = sc.materials[beer - 1];
			W = W * mkc(fExp(-hh.x * bm.beer_sigma_a[0]), fExp(-hh.x * bm.beer_sigma_a[1]), fExp(-hh.x * bm.beer_sigma_a[2]));
		}
		const SlotInfo si = slotInfo(wp, pixels, slot);
		const int ray_division = stDivision(packed);

		if((KIND == 2 ? stKind(packed) : KIND) == TASK_INTEGRATE)
		{
			// ---- PathIntegrator::integrate / DirectLightIntegrator::integrate ----
			if(MT == MT_MISS || tri < 0)
			{
				if(sc.bg.type != YAFGPU_BG_NONE && !ig.transp_refr_bg) accAdd(acc, slot, W * bgEval(sc.bg, dir));   // integrator_path_tracer.cc:328
				if(packed & YG_ST_ALPHA)
				{	// the refraction chain left the scene: this integrate() returns its initial alpha (integrator_path_tracer.cc:122-123)
					const float4 ab = in.thr[i];
					((float *)(acc + slot))[3] = alphaEncode(ab.x + ab.y * (ig.transp_bg ? 0.f : 1.f));
				}
				else if(ig.transp_bg && !ig.transp_refr_bg && stRaylevel(packed) == 0) ((float *)(acc + slot))[3] = alphaEncode(0.f);   // primary ray missed: alpha stays at its initial 0 (integrator_path_tracer.cc:122)
				continue;
			}
			SurfPt sp;
			getSurface(sc, tri, hh.y, hh.z, from + hh.x * dir, sp); // Point3 h = ray.from_ + z * ray.dir_ (scene.cc:908)
			const unsigned mono = packed & YG_ST_MONO;
			sp.chromatic = mono ? 0 : 1; sp.wavelength = mono ? __uint_as_float(st.y) : 0.f;

			// materials with shader nodes (kernel class YG_MAT_NODES): the node pass of initBsdf() resolves the bound shaders into a
			// thread-local copy of the record (nodes.cuh); every other kernel reads the record in place
			const MatDev *mp = &sc.materials[sp.mat];
			MatDev ml;
			if(MT == MT_NODES) { ml = *mp; evalMaterialNodes(sc, ml, sp, tri, hh.y, hh.z); mp = &ml; }
			const MatDev &m = *mp;
			matInitBsdf<MT>(m, sp);
			const unsigned bsdfs = m.bsdf_flags;
			const V3 wo = -dir;
			const int raylevel = stRaylevel(packed);
			const int additional_depth = max(stAddDepth(packed), m.additional_depth);
			const float2 dcv = in.dc[i];
			// "bg_transp_refract": alpha of the camera sample = A + B * (m + (1 - m) * alpha_child); carried on by the last refracted child, settled here when there is none
			float alpha_a = 0.f, alpha_b = 0.f; int alpha_child = -1;
			const bool alpha_chain = (packed & YG_ST_ALPHA) != 0;
			if(alpha_chain)
			{
				const float4 ab = in.thr[i];
				const float m_alpha = matAlpha<MT>(sc, m, sp, wo);
				alpha_a = ab.x + ab.y * m_alpha; alpha_b = ab.y * (1.f - m_alpha);
			}
			if(bsdfs & BsdfEmit) accAdd(acc, slot, W * matEmit<MT>(sc, m, sp, wo, !(packed & YG_ST_NOLIGHTS)));
			if(bsdfs & BsdfDiffuse)
			{	// estimateAllDirectLight integrator_montecarlo.cc:47-60
				for(int l = 0; l < sc.n_lights; ++l) doLightEstimation<MT>(cx, m, slot, si, l, sp, wo, (unsigned)l, ray_division, W);
				if(ig.type == YAFGPU_INT_DIRECTLIGHT && ig.do_ao)
				{	// MonteCarloIntegrator::sampleAmbientOcclusion integrator_montecarlo.cc:1030-1088 (only DirectLight calls it, :145)
					int n = ig.ao_samples;
					if(ray_division > 1) n = max(1, n / ray_division);
					const unsigned offs = (unsigned)n * (unsigned)si.pixel_sample + si.sampling_offs;
					Halton2 hal_2; HaltonB<3> hal_3;
					hal_2.setStart(offs - 1u); hal_3.setStart(offs - 1u);
					const float ao_tmin = ig.shadow_bias_auto ? ig.shadow_bias * fmaxf(1.f, length(sp.p)) : ig.shadow_bias;
					const Col ao_col = mkc(ig.ao_col[0], ig.ao_col[1], ig.ao_col[2]);
					for(int k = 0; k < n; ++k)
					{
						float s_1 = hal_2.getNext(), s_2 = hal_3.getNext();
						if(ray_division > 1) { s_1 = addMod1(s_1, dcv.x); s_2 = addMod1(s_2, dcv.y); }
						BSample s; s.s1 = s_1; s.s2 = s_2; s.flags = BsdfGlossy | BsdfDiffuse | BsdfReflect; s.pdf = 0.f; s.sampled = BsdfNone;
						float w = 0.f; V3 adir = mk(0.f);
						const Col surf_col = matSample<MT>(sc, m, sp, wo, adir, s, w);
						if((MT == MT_NODES ? m.obj_flags : m.bsdf_flags) & BsdfEmit) accAdd(acc, slot, W * ((matEmit<MT>(sc, m, sp, wo, true) * s.pdf) / (float)n));
						const float cosv = fabsf(dot(sp.n, adir));
						pushShadow(cx, slot, sp.p, ao_tmin, adir, ig.ao_dist, W * ((((ao_col * surf_col) * cosv) * w) / (float)n));
					}
				}
			}
			if(ig.type == YAFGPU_INT_PATHTRACING)
			{	// integrator_path_tracer.cc:174-298: first path segment, sampled here; the rest of each path is a TASK_PATH chain
				unsigned path_flags = ig.no_recursive ? BsdfAll : BsdfDiffuse;
				if(bsdfs & path_flags)
				{
					path_flags |= (BsdfDiffuse | BsdfReflect | BsdfTransmit);
					const int n_samples = max(1, ig.n_paths / ray_division);
					const float inv_n = 1.f / (float)n_samples;
					for(int k = 0; k < n_samples; ++k)
					{
						const unsigned offs = (unsigned)ig.n_paths * (unsigned)si.pixel_sample + si.sampling_offs + (unsigned)k;
						if(sp.chromatic) sp.wavelength = riS(offs);   // integrator_path_tracer.cc:196-197
						float s_1 = riVdC(offs);
						float s_2 = (float)scrHaltonT(*sc.faure, 2, offs);
						if(ray_division > 1) { s_1 = addMod1(s_1, dcv.x); s_2 = addMod1(s_2, dcv.y); }
						BSample s; s.s1 = s_1; s.s2 = s_2; s.flags = path_flags; s.pdf = 0.f; s.sampled = BsdfNone;
						float w = 0.f; V3 wi = mk(0.f);
						Col scol = matSample<MT>(sc, m, sp, wo, wi, s, w);
						scol = scol * w;
						if(isBlack(scol) || s.sampled == BsdfNone) continue; // contributes exactly zero in the reference as well
						pushTask(cx, slot, sp.p, ig.ray_min_dist, wi, (W * scol) * inv_n, scol, offs,
						         packState(TASK_PATH, raylevel, 0, ray_division, 0, 0) | mono, 0u, 0.f, 0.f);
					}
				}
			}
			// ---- recursiveRaytrace integrator_montecarlo.cc:782-1028 ----
			const int rl = raylevel + 1;
			if(rl <= ig.r_depth + additional_depth)
			{
				const unsigned child_wl = mono ? st.y : 0u;   // a single-wavelength task hands its wavelength down
				if((bsdfs & BsdfDispersive) && sp.chromatic)
				{	// dispersive branch, integrator_montecarlo.cc:796-858: 8 wavelengths (fewer below a ray division), each refracted with its own IOR
					// and traced as a single-wavelength integrate() whose result is tinted by the wavelength's colour
					int dsam = 8;
					const int old_division = ray_division, old_offset = (int)st.w;
					if(old_division > 1) dsam = max(1, dsam / old_division);
					const int new_division = old_division * dsam;
					int branch = new_division * old_offset;
					const float d_1 = 1.f / (float)dsam;
					const float ss_1 = riS((uint32_t)si.pixel_sample + si.sampling_offs);
					for(int ns = 0; ns < dsam; ++ns)
					{
						float wl = ((float)ns + ss_1) * d_1;
						const float dc1 = (float)scrHaltonT(*sc.faure, 2 * rl + 1, (unsigned)branch + si.sampling_offs);
						const float dc2 = (float)scrHaltonT(*sc.faure, 2 * rl + 2, (unsigned)branch + si.sampling_offs);
						if(old_division > 1) wl = addMod1(wl, dcv.x);
						const unsigned ray_offset = (unsigned)branch;
						++branch;
						sp.wavelength = wl;   // stays set after the loop: getSpecular below reads the LAST sample's wavelength, as the reference's state does
						BSample s; s.s1 = 0.5f; s.s2 = 0.5f; s.flags = BsdfReflect | BsdfTransmit | BsdfDispersive; s.pdf = 0.f; s.sampled = BsdfNone;
						float w = 0.f; V3 wi = mk(0.f);
						const Col mcol = matSample<MT>(sc, m, sp, wo, wi, s, w);
						if(s.pdf > 1.0e-6f && (s.sampled & BsdfDispersive))
						{
							const Col f = ((mcol * wl2Rgb(sc, wl)) * w) * d_1;
							if(!isBlack(f))
								pushTask(cx, slot, sp.p, ig.ray_min_dist, wi, W * f, mkc(1.f), __float_as_uint(wl),
								         packState(TASK_INTEGRATE, rl, 0, new_division, 0, additional_depth) | YG_ST_MONO | YG_ST_NOLIGHTS, ray_offset, dc1, dc2);
						}
					}
				}
				if((bsdfs & BsdfGlossy) && rl < 20)
				{
					int gsam = 8;
					const int old_division = ray_division, old_offset = (int)st.w;
					if(old_division > 1) gsam = max(1, gsam / old_division);
					const int new_division = old_division * gsam;
					int branch = new_division * old_offset;
					const unsigned offs = (unsigned)gsam * (unsigned)si.pixel_sample + si.sampling_offs;
					const float d_1 = 1.f / (float)gsam;
					Halton2 hal_2; HaltonB<3> hal_3;
					hal_2.setStart(offs); hal_3.setStart(offs);
					for(int ns = 0; ns < gsam; ++ns)
					{
						const float dc1 = (float)scrHaltonT(*sc.faure, 2 * rl + 1, (unsigned)branch + si.sampling_offs);
						const float dc2 = (float)scrHaltonT(*sc.faure, 2 * rl + 2, (unsigned)branch + si.sampling_offs);
						const unsigned ray_offset = (unsigned)branch;
						++branch;
						const float s_1 = hal_2.getNext(), s_2 = hal_3.getNext();
						const unsigned child = packState(TASK_INTEGRATE, rl, 0, new_division, 0, additional_depth) | mono;
						// material->getFlags() here, not the flags initBsdf returned (integrator_montecarlo.cc:895-919): they differ under a mask_mat
						const unsigned oflags = MT == MT_NODES ? m.obj_flags : m.bsdf_flags;
						if(oflags & BsdfGlossy)
						{
							if((oflags & BsdfReflect) && !(oflags & BsdfTransmit))
							{
								BSample s; s.s1 = s_1; s.s2 = s_2; s.flags = BsdfGlossy | BsdfReflect; s.pdf = 0.f; s.sampled = BsdfNone;
								float w = 0.f; V3 wi = mk(0.f);
								const Col mcol = matSample<MT>(sc, m, sp, wo, wi, s, w);
								const Col f = (mcol * w) * d_1;
								if(!isBlack(f)) pushTask(cx, slot, sp.p, ig.ray_min_dist, wi, W * f, mkc(1.f), child_wl, child, ray_offset, dc1, dc2, beerOf(m, sp, wi));
							}
							else if(MT == MT_NODES && m.is_mask) {} // MaskMaterial inherits the two-lobe Material::sample, which returns black and samples nothing
							else if((oflags & BsdfReflect) && (oflags & BsdfTransmit))
							{
								BSample s; s.s1 = s_1; s.s2 = s_2; s.flags = BsdfGlossy | BsdfAllGlossy; s.pdf = 0.f; s.sampled = BsdfNone;
								Col mcol[2] = {mkc(0.f), mkc(0.f)}; float w[2] = {0.f, 0.f}; V3 d2[2] = {mk(0.f), mk(0.f)};
								mcol[0] = matSample2<MT>(sc, m, sp, wo, d2, mcol[1], s, w);
								if((s.sampled & BsdfReflect) && !(s.sampled & BsdfDispersive))
								{
									const Col f = (mcol[0] * w[0]) * d_1;
									if(!isBlack(f)) pushTask(cx, slot, sp.p, ig.ray_min_dist, d2[0], W * f, mkc(1.f), child_wl, child, ray_offset, dc1, dc2, beerOf(m, sp, d2[0]));
								}
								if(s.sampled & BsdfTransmit)
								{	// `alpha = integ.a_` (:954): the last of these samples decides, and its ray is traced even when it carries no colour
									const Col f = (mcol[1] * w[1]) * d_1;
									if(!isBlack(f) || alpha_chain)
									{
										const int k = pushTask(cx, slot, sp.p, ig.ray_min_dist, d2[1], W * f, mkc(1.f), child_wl, child, ray_offset, dc1, dc2, beerOf(m, sp, d2[1]));
										if(alpha_chain) alpha_child = k;
									}
								}
							}
						}
					}
				}
				if((bsdfs & (BsdfSpecular | BsdfFilter)) && rl < 20)
				{
					bool refl = false, refr = false;
					V3 d2[2]; Col rcol[2];
					matSpecular<MT>(sc, m, sp, wo, rl, refl, refr, d2, rcol);
					const unsigned child = packState(TASK_INTEGRATE, rl, 0, ray_division, 0, additional_depth) | mono;
					if(refl && !isBlack(rcol[0])) pushTask(cx, slot, sp.p, ig.ray_min_dist, d2[0], W * rcol[0], mkc(1.f), child_wl, child, st.w, dcv.x, dcv.y, beerOf(m, sp, d2[0]));
					if(refr && (!isBlack(rcol[1]) || alpha_chain))
					{
						V3 o = sp.p;
						float tb = m.transp_bias_factor;
						if(tb > 0.f)
						{
							if(m.transp_bias_mult) tb *= (float)rl;
							o = sp.p + d2[1] * tb;
						}
						const int k = pushTask(cx, slot, o, ig.ray_min_dist, d2[1], W * rcol[1], mkc(1.f), child_wl, child, st.w, dcv.x, dcv.y, beerOf(m, sp, d2[1]));
						if(alpha_chain) alpha_child = k;   // `alpha = integ.a_` (:1023) overrides the glossy samples'
					}
				}
			}
			if(alpha_chain)
			{
				if(alpha_child >= 0)
				{
					cx.out.st[alpha_child].z |= YG_ST_ALPHA;
					cx.out.thr[alpha_child] = make_float4(alpha_a, alpha_b, 1.f, 0.f);
				}
				else ((float *)(acc + slot))[3] = alphaEncode(alpha_a + alpha_b * (ig.transp_bg ? 0.f : 1.f));
			}
		}
		else
		{
			// ---- one vertex of the path loop, integrator_path_tracer.cc:214-294 ----
			const int depth = stDepth(packed);
			const bool caustic = stCaustic(packed);
			if(MT == MT_MISS || tri < 0)
			{
				if(depth >= 1 && caustic && sc.bg.type != YAFGPU_BG_NONE && sc.bg.has_ibl && sc.bg.shoots_caustic) accAdd(acc, slot, W * bgEval(sc.bg, dir));
				continue;
			}
			SurfPt sp;
			getSurface(sc, tri, hh.y, hh.z, from + hh.x * dir, sp);
			const unsigned mono = packed & YG_ST_MONO;
			sp.chromatic = mono ? 0 : 1; sp.wavelength = mono ? 0.f : riS(st.y);   // the path keeps the wavelength of its first segment (integrator_path_tracer.cc:196-197)
			// materials with shader nodes (kernel class YG_MAT_NODES): the node pass of initBsdf() resolves the bound shaders into a
			// thread-local copy of the record (nodes.cuh); every other kernel reads the record in place
			const MatDev *mp = &sc.materials[sp.mat];
			MatDev ml;
			if(MT == MT_NODES) { ml = *mp; evalMaterialNodes(sc, ml, sp, tri, hh.y, hh.z); mp = &ml; }
			const MatDev &m = *mp;
			matInitBsdf<MT>(m, sp);
			const unsigned bsdfs = m.bsdf_flags;
			const V3 pwo = -dir;
			const float4 th = in.thr[i];
			Col thr = mkc(th.x, th.y, th.z);
			Col Wp = W;
			const unsigned offs = st.y;
			bool direct = (depth == 0) || (bsdfs & BsdfDiffuse);
			if(YG_CAN_ABSORB(MT) && depth >= 1 && m.absorb && dot(sp.n, pwo) < 0.f)
			{	// integrator_path_tracer.cc:276-279: arriving from inside an absorbing medium, before the roulette
				const Col vcol = mkc(fExp(-hh.x * m.beer_sigma_a[0]), fExp(-hh.x * m.beer_sigma_a[1]), fExp(-hh.x * m.beer_sigma_a[2]));
				thr = thr * vcol; Wp = Wp * vcol;
			}
			if(depth >= 1 && depth > ig.rr_min_bounces)
			{	// Russian roulette :282-288
				const float random_value = hashRand(wp.seed ^ ((unsigned)si.x | ((unsigned)si.y << 16)), offs, (unsigned)depth);   // (pass, pixel | sample, path via offs | depth): independent of wave and partition
				const float probability = maxc(thr);
				if(probability <= 0.f || probability < random_value) continue;
				const float ip = 1.f / probability;
				thr = thr * ip; Wp = Wp * ip;
			}
			if(direct && sc.n_lights > 0 && (bsdfs & (BsdfDiffuse | BsdfGlossy)))
			{	// estimateOneDirectLight integrator_montecarlo.cc:62-76 (a material without diffuse/glossy lobes evaluates to black)
				const float u = hashRand((wp.seed + 0x51ED270Bu) ^ ((unsigned)si.x | ((unsigned)si.y << 16)), offs, (unsigned)depth);
				const int lnum = min((int)(u * (float)sc.n_lights), sc.n_lights - 1);
				doLightEstimation<MT>(cx, m, slot, si, lnum, sp, pwo, (unsigned)lnum, ray_division, Wp * (float)sc.n_lights);
			}
			if(bsdfs & BsdfEmit)
			{
				if(depth == 0) accAdd(acc, slot, Wp * matEmit<MT>(sc, m, sp, pwo, false));  // include_lights_ = false (:212)
				else if(caustic) accAdd(acc, slot, Wp * matEmit<MT>(sc, m, sp, pwo, true)); // include_lights_ = caustic (:253,290)
			}
			const int nd = depth + 1;
			if(nd < ig.max_bounces)
			{
				BSample s; s.flags = BsdfAll; s.pdf = 0.f; s.sampled = BsdfNone;
				s.s1 = (float)scrHaltonT(*sc.faure, 4 * nd + 3, offs);
				s.s2 = (float)scrHaltonT(*sc.faure, 4 * nd + 4, offs);
				float w = 0.f; V3 wi = mk(0.f);
				Col scol = matSample<MT>(sc, m, sp, pwo, wi, s, w);
				scol = scol * w;
				if(isBlack(scol)) continue;
				const bool ncaustic = ig.trace_caustics && (s.sampled & (BsdfSpecular | BsdfGlossy | BsdfFilter));
				pushTask(cx, slot, sp.p, ig.ray_min_dist, wi, Wp * scol, thr * scol, offs,
				         packState(TASK_PATH, stRaylevel(packed), nd, ray_division, ncaustic ? 1 : 0, 0) | mono, 0u, 0.f, 0.f);
			}
		}
	}
}

// ---- K4: shadow rays ---------------------------------------------------------------------------------------------------
// TriKdTree::intersectTs kdtree_triangle.cc:983-1162: every shadow-visible triangle in (tmin, dist); opaque -> blocked;
// transparent -> filt *= getTransparency, at most max_depth of them (a BVH holds each triangle once, so the reference's
// std::set of already-filtered triangles is implicit).
struct TranspF
{
	const SceneDev *sc; V3 o, d; float tmin; int max_depth, depth; Col filt; bool blocked;
	__device__ __forceinline__ bool operator()(const float4 e1i, const float4 e2f, float t, float u, float v, float &z)
	{
		const int fl = __float_as_int(e2f.w);
		if(!(t < z && t >= tmin && (fl & TRI_VIS_SHADOW))) return false;
		if(!(fl & TRI_TRANSPARENT) || depth >= max_depth) { blocked = true; return true; }
		SurfPt sp;
		getSurface(*sc, __float_as_int(e1i.w), u, v, o + t * d, sp);
		filt = filt * matTransparency(*sc, sc->materials[sp.mat], sp, d);
		++depth;
		return false;
	}
};

template<int NS>
__device__ bool traceShadowTransp(const SceneDev &sc, V3 o, V3 d, float tmin, float dist, int max_depth, Col &filt, TravStack<NS> &st)
{
	TranspF f; f.sc = &sc; f.o = o; f.d = d; f.tmin = tmin; f.max_depth = max_depth; f.depth = 0; f.filt = mkc(1.f); f.blocked = false;
	float z = dist;
	traverseWide<false, false, NS>(sc.accel, o, d, 0.f, z, st, f, nullptr);
	filt = f.filt;
	return f.blocked;
}

struct ShadowIO
{
	ShadowQueue sq; float4 *acc;
	__device__ __forceinline__ void load(int i, V3 &o, V3 &d, float &tmin, float &dist) const
	{
		const float4 a = sq.ray_o[i], b = sq.ray_d[i];
		d = mk(b.x, b.y, b.z);
		// Scene::isShadowed scene.cc:965-970: shift the origin by tmin, shorten the segment by 2 tmin
		o = mk(a.x, a.y, a.z) + d * a.w; tmin = a.w;
		dist = (b.w < 0.f) ? __int_as_float(0x7f800000) : b.w - 2 * a.w;
	}
	__device__ __forceinline__ void store(int i, const HitD &h, unsigned) const
	{
		if(h.tri >= 0) return; // blocked
		const float4 c = sq.contrib[i];
		accAdd(acc, __float_as_int(c.w), mkc(c.x, c.y, c.z));
	}
};

// opaque shadows: any-hit traversal (REGEN as in k_trace); transparent shadows (transpShad): filter accumulation along the segment
template<bool COUNT, bool REGEN>
__global__ void __launch_bounds__(128) k_shadow(const __grid_constant__ SceneDev sc, ShadowQueue sq, const int *__restrict__ n_shadow, int shadow_cap, float4 *__restrict__ acc,
                                                int *__restrict__ next, int refill_below, unsigned long long *__restrict__ counters, const WaveCtl *__restrict__ ctl)
{
	if(ctl->overflow) return;
	const int n = min(*n_shadow, shadow_cap);
	TravCounters tc; tc.nodes = 0; tc.tris = 0;
	__shared__ uint2 s_stack[YG_STACK_SHARED * 128];
	TravStack<YG_STACK_SHARED> st; bindSharedStack(st, s_stack);
	if(REGEN && !sc.integ.tr_shad)
	{
		ShadowIO io; io.sq = sq; io.acc = acc;
		traversePersistent<true, COUNT>(sc.accel, n, next, io, st, &tc, refill_below);
	}
	else
	for(int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
	{
		const float4 a = sq.ray_o[i], b = sq.ray_d[i], c = sq.contrib[i];
		const V3 dir = mk(b.x, b.y, b.z);
		// Scene::isShadowed scene.cc:965-970: shift the origin by tmin, shorten the segment by 2 tmin
		const V3 o = mk(a.x, a.y, a.z) + dir * a.w;
		const float dist = (b.w < 0.f) ? __int_as_float(0x7f800000) : b.w - 2 * a.w;
		Col contrib = mkc(c.x, c.y, c.z);
		if(sc.integ.tr_shad)
		{
			Col filt;
			if(traceShadowTransp(sc, o, dir, a.w, dist, sc.integ.s_depth, filt, st)) continue;
			contrib = contrib * filt;
		}
		else if(traceShadowS<COUNT>(sc.accel, o, dir, dist, st, &tc) >= 0) continue;
		accAdd(acc, __float_as_int(c.w), contrib);
	}
	if(COUNT) { atomicAdd(&counters[CNT_NODES_SHADOW], tc.nodes); atomicAdd(&counters[CNT_TRIS_SHADOW], tc.tris); }
}

// ---- K9: ImageFilm::addSample imagefilm.cc:925-1015 (Combined pass) --------------------------------------------------------
__device__ __forceinline__ int round2Int(double v) { return (int)(v + (.5 - 1.4e-11)); } // util_math.h:34-43 (FAST_INT off)

__global__ void __launch_bounds__(256) k_film_add(FilmDev film, WaveParams wp, const uint32_t *__restrict__ pixels, const float4 *__restrict__ acc, const WaveCtl *__restrict__ ctl,
                                                  int keep_nonfinite, unsigned long long *__restrict__ counters, int it_end)
{
	if(ctl->overflow) return; // the wave is discarded and re-rendered in smaller pieces
	if(ctl->n_tasks[it_end] != 0) return; // tasks are still queued behind the iterations launched so far: the host launches more and splats afterwards
	for(int slot = blockIdx.x * blockDim.x + threadIdx.x; slot < wp.n_slots; slot += gridDim.x * blockDim.x)
	{
		const SlotInfo si = slotInfo(wp, pixels, slot);
		float dx, dy;
		samplePos(wp, si, dx, dy);
		const float4 a = acc[slot];
		// A path can return NaN/Inf where the reference's own arithmetic does (e.g. RoughGlassMaterial::sample, material_rough_glass.cc:225: |wo_h wi_h / (wi_n wo_n)| * G
		// with wo_n == 0 and G == 0 is inf * 0; about one camera sample in 10^9 of the bench scene).  The reference splats it and the pixel stays NaN for good; here the
		// sample is left out (colour AND weight) and counted in yafgpu_stats.nonfinite_samples, unless option "keep_nonfinite_samples" asks for the reference's behaviour.
		if(!keep_nonfinite && !(isfinite(a.x) && isfinite(a.y) && isfinite(a.z))) { atomicAdd(&counters[CNT_NONFINITE], 1ull); continue; }
		Col col = clampProportional(mkc(a.x, a.y, a.z), film.clamp_samples);
		const float alpha = alphaEncode(a.w);   // decode == encode (XOR); 0 reads as 1
		const int x = si.x, y = si.y;
		const double fw = (double)film.filterw;
		const int dx_0 = max(film.cx0 - x, round2Int((double)dx - fw));
		const int dx_1 = min(film.cx1 - x - 1, round2Int((double)dx + fw - 1.0));
		const int dy_0 = max(film.cy0 - y, round2Int((double)dy - fw));
		const int dy_1 = min(film.cy1 - y - 1, round2Int((double)dy + fw - 1.0));
		const double x_offs = (double)dx - 0.5, y_offs = (double)dy - 0.5;
		int x_index[9], y_index[9];
		for(int i = dx_0, n = 0; i <= dx_1 && n < 9; ++i, ++n) x_index[n] = (int)floor(fabs(((double)i - x_offs) * (double)film.table_scale));
		for(int i = dy_0, n = 0; i <= dy_1 && n < 9; ++i, ++n) y_index[n] = (int)floor(fabs(((double)i - y_offs) * (double)film.table_scale));
		const int x_0 = x + dx_0, x_1 = x + dx_1, y_0 = y + dy_0, y_1 = y + dy_1;
		for(int j = y_0; j <= y_1; ++j)
			for(int i = x_0; i <= x_1; ++i)
			{
				const float filter_wt = film.table[y_index[j - y_0] * 16 + x_index[i - x_0]];
				float *p = film.pix + 5 * ((size_t)(j - film.cy0) * film.w + (i - film.cx0));
				atomicAdd(p + 0, col.r * filter_wt);
				atomicAdd(p + 1, col.g * filter_wt);
				atomicAdd(p + 2, col.b * filter_wt);
				atomicAdd(p + 3, alpha * filter_wt);
				atomicAdd(p + 4, filter_wt);
			}
	}
}

// ---- host side ---------------------------------------------------------------------------------------------------------
static unsigned hashSeed(unsigned a, unsigned b, unsigned c)
{
	unsigned h = a * 0x9E3779B1u + 0x7F4A7C15u;
	h ^= b + (h << 6) + (h >> 2); h *= 0x85EBCA6Bu; h ^= h >> 13;
	h ^= c + (h << 6) + (h >> 2); h *= 0xC2B2AE35u; h ^= h >> 16;
	return h;
}
void freeWavefront(yafgpu_ctx *c)
{
	Wavefront &w = c->wf;
	for(int k = 0; k < 2; ++k)
	{
		cudaFree(w.q[k].ray_o); cudaFree(w.q[k].ray_d); cudaFree(w.q[k].hit); cudaFree(w.q[k].wgt); cudaFree(w.q[k].thr); cudaFree(w.q[k].st); cudaFree(w.q[k].dc);
		w.q[k] = TaskQueue();
	}
	cudaFree(w.sq.ray_o); cudaFree(w.sq.ray_d); cudaFree(w.sq.contrib); w.sq = ShadowQueue();
	cudaFree(w.acc); cudaFree(w.pixels); cudaFree(w.ctl); cudaFree(w.key); cudaFree(w.perm); cudaFree(w.hist);
	w.key = w.perm = w.hist = nullptr;
	if(w.h_ctl) cudaFreeHost(w.h_ctl);
	w.acc = nullptr; w.pixels = nullptr; w.ctl = nullptr; w.h_ctl = nullptr; w.pixels_cap = 0;
	w.allocated = false; w.pixels_key.clear();
}

static cudaError_t allocWavefront(yafgpu_ctx *c, int wave_cap, int task_cap, int shadow_cap)
{
	Wavefront &w = c->wf;
	if(w.allocated && w.wave_cap == wave_cap && w.task_cap == task_cap && w.shadow_cap == shadow_cap) return cudaSuccess;
	uint32_t *pix = w.pixels; size_t pcap = w.pixels_cap; std::vector<int> key = w.pixels_key; // keep the pixel list
	w.pixels = nullptr;
	freeWavefront(c);
	w.pixels = pix; w.pixels_cap = pcap; w.pixels_key = key;
	cudaError_t e = cudaSuccess;
#define A(p, T, n) if(e == cudaSuccess) e = cudaMalloc(&(p), sizeof(T) * (size_t)(n))
	for(int k = 0; k < 2; ++k)
	{
		A(w.q[k].ray_o, float4, task_cap); A(w.q[k].ray_d, float4, task_cap); A(w.q[k].hit, float4, task_cap); A(w.q[k].wgt, float4, task_cap);
		A(w.q[k].thr, float4, task_cap); A(w.q[k].st, uint4, task_cap); A(w.q[k].dc, float2, task_cap);
	}
	A(w.sq.ray_o, float4, shadow_cap); A(w.sq.ray_d, float4, shadow_cap); A(w.sq.contrib, float4, shadow_cap);
	A(w.acc, float4, wave_cap); A(w.ctl, WaveCtl, 1);
	A(w.key, int, task_cap); A(w.perm, int, task_cap); A(w.hist, int, 2 * YG_MAX_BUCKETS);
#undef A
	if(e == cudaSuccess) e = cudaMallocHost(&w.h_ctl, sizeof(WaveCtl));
	if(e != cudaSuccess) { uint32_t *p2 = w.pixels; w.pixels = nullptr; freeWavefront(c); w.pixels = p2; return e; }
	w.wave_cap = wave_cap; w.task_cap = task_cap; w.shadow_cap = shadow_cap; w.allocated = true;
	return cudaSuccess;
}

// Owner of tile (tx, ty) among `n` partitions: (tx + ty * stride) mod n with stride = the integer nearest to 0.618 n that is coprime with n.  The plain raster
// rule (t mod n) gives a partition whole columns of tiles whenever the row length shares a factor with n (60 tiles per row, 8 GPUs: ranks k and k + 4 own the same
// columns), and the cost of a tile varies mostly across the frame: the slowest partition of the bench frame was 6 % above the mean.  The shifted rule walks every
// partition through all columns within n rows (tools/partition_balance.py).  Same rule in libyafaray_b200/parallel.py owner_map.
#ifndef YG_PARTITION_UNIT
#define YG_PARTITION_UNIT 8
#endif
static int partitionStride(int n)
{
	auto gcd = [](int a, int b) { while(b) { const int t = a % b; a = b; b = t; } return a; };
	int s = std::max(1, (int)floor(0.618 * n + 0.5));
	while(gcd(s, n) != 1) ++s;
	return s % n ? s % n : 1;
}

// pixel list of a pass in tile order: ImageSplitter raster tiles (imagesplitter.cc:37-50), rows inside a tile
static int buildPixelList(yafgpu_ctx *c, const yafgpu_pass_desc *p, size_t &n_out)
{
	Wavefront &w = c->wf;
	const FilmDev &f = c->film;
	const int part_count = p->part_count > 1 ? p->part_count : 1, part_index = p->part_count > 1 ? p->part_index : 0;
	// a partitioned pass interleaves blocks of at most 8 x 8 pixels: the finer the unit, the closer the partitions' costs (bench frame, 8 parts: slowest partition
	// 3.8 % above the mean with 32 x 32 tiles, 2.2 % with 8 x 8; tools/partition_balance.py); an unpartitioned pass walks the reference's tile_size grid
	const int ts = part_count > 1 ? std::min(c->tile_size, YG_PARTITION_UNIT) : c->tile_size;
	const bool masked = p->adaptive && p->resample_mask;
	std::vector<int> key = {f.w, f.h, f.cx0, f.cy0, ts, part_index, part_count};
	if(!masked && w.pixels && w.pixels_key == key) { n_out = w.pixels_n; return YAFGPU_OK; }
	std::vector<uint32_t> px;
	px.reserve((size_t)f.w * f.h / part_count + 1024);
	const int ntx = (f.w + ts - 1) / ts, nty = (f.h + ts - 1) / ts;
	const int pstride = partitionStride(part_count);
	for(int ty = 0; ty < nty; ++ty) for(int tx = 0; tx < ntx; ++tx)
	{
		if(part_count > 1 && (tx + ty * pstride) % part_count != part_index) continue;
		const int x0 = tx * ts, y0 = ty * ts, x1 = std::min(x0 + ts, f.w), y1 = std::min(y0 + ts, f.h);
		for(int y = y0; y < y1; ++y) for(int x = x0; x < x1; ++x)
		{
			if(masked && !p->resample_mask[(size_t)y * f.w + x]) continue;
			px.push_back((uint32_t)(x + f.cx0) | ((uint32_t)(y + f.cy0) << 16));
		}
	}
	if(px.size() > w.pixels_cap)
	{
		cudaFree(w.pixels); w.pixels = nullptr; w.pixels_cap = 0;
		YG_CUDA(c, cudaMalloc(&w.pixels, sizeof(uint32_t) * (px.size() + 1)));
		w.pixels_cap = px.size() + 1;
	}
	if(!px.empty()) YG_CUDA(c, cudaMemcpyAsync(w.pixels, px.data(), sizeof(uint32_t) * px.size(), cudaMemcpyHostToDevice, c->stream));
	YG_CUDA(c, cudaStreamSynchronize(c->stream)); // px goes out of scope
	w.pixels_n = px.size(); n_out = px.size();
	if(masked) w.pixels_key.clear(); else w.pixels_key = key;
	return YAFGPU_OK;
}

static int runWave(yafgpu_ctx *c, const WaveParams &wp, int max_iters, bool &overflow)
{
	Wavefront &w = c->wf;
	cudaStream_t st = c->stream;
	const bool count = (c->flags & YAFGPU_FLAG_COUNT_TRAVERSAL) != 0;
	// Default schedule: both traversal kernels regenerate rays on scenes whose rays are long walks (10^7 triangles: shadow kernel 10.4 -> 7.1 ms per 2 spp); on a scene
	// of a few dozen triangles a shadow ray is a handful of node visits and the refill atomics cost more than the idle lanes did (Cornell: 3.0 -> 5.4 ms per step,
	// tools/cornell_regen_ab.py), so there only the closest-hit kernel regenerates.
	const int rr = c->ray_regeneration >= 0 ? c->ray_regeneration : (YG_DEFAULT_REGEN == 1 && c->scene.accel.n_tris < (1 << 15) ? 2 : YG_DEFAULT_REGEN);
	const bool regen_t = rr == 1 || rr == 2, regen_s = rr == 1 || rr == 3;
	const int refill = c->regen_threshold;
	int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
	YG_CUDA(c, cudaMemsetAsync(w.ctl, 0, sizeof(WaveCtl), st));
	YG_CUDA(c, cudaMemsetAsync(w.acc, 0, sizeof(float4) * (size_t)wp.n_slots, st));
	YG_CUDA(c, cudaMemcpyAsync(&w.ctl->n_tasks[0], &wp.n_slots, sizeof(int), cudaMemcpyHostToDevice, st));
	const int g256 = std::min((wp.n_slots + 255) / 256, sms * 8);
	k_raygen<<<g256, 256, 0, st>>>(c->scene, wp, w.pixels, w.q[0]);
	const int gt = sms * 16, gs = sms * 8;
	const bool timed = (c->flags & YAFGPU_FLAG_TIME_STAGES) != 0;
	std::vector<cudaEvent_t> &ev = c->stage_events;
	size_t ne = 0;
	auto mark = [&]() { if(!timed) return; if(ne >= ev.size()) { cudaEvent_t e; cudaEventCreate(&e); ev.push_back(e); } cudaEventRecord(ev[ne++], st); };
	mark();
	// The host cannot see the queue lengths without a round trip, so it launches the number of iterations the previous wave of this scene needed (+ 1) and lets the
	// DEVICE decide whether that was enough: k_film_add only splats when nothing is queued behind the last launched iteration.  If tasks remain (seen in the control
	// block the wave-end synchronisation reads anyway), four more iterations and another k_film_add follow.  Empty tail iterations launch nothing any more.
	int it_end = (c->iters_hint > 0 && c->adaptive_iters) ? std::min(max_iters, c->iters_hint) : max_iters;
	int it = 0;
more_iterations:
	for(; it < it_end; ++it)
	{
		int n_launch = 0;
		TaskQueue &in = w.q[it & 1], &out = w.q[(it + 1) & 1];
		SortBins bins; bins.tri_mat = c->d_tri_mat; bins.mat_bucket = c->d_mat_bucket; bins.n_mat = c->n_mat_buckets; bins.key = w.key; bins.hist = w.hist;
		const int nb = 2 * c->n_mat_buckets + 1;
		YG_CUDA(c, cudaMemsetAsync(w.hist, 0, sizeof(int) * nb, st));
		const size_t hist_smem = nb <= YG_SHARED_BUCKETS ? sizeof(int) * nb : 0;
#define YG_TRACE(C, R) k_trace<C, R><<<gt, 128, hist_smem, st>>>(c->scene.accel, in, &w.ctl->n_tasks[it], wp.task_cap, bins, &w.ctl->fetch_trace[it], refill, c->d_counters, w.ctl)
		if(count) { if(regen_t) YG_TRACE(true, true); else YG_TRACE(true, false); } else { if(regen_t) YG_TRACE(false, true); else YG_TRACE(false, false); }
#undef YG_TRACE
		mark();
		k_bin_scan<<<1, 1024, 0, st>>>(w.hist, w.hist + YG_MAX_BUCKETS, nb);
		if(nb <= YG_SHARED_BUCKETS) k_bin_scatter_block<<<sms * 4, 256, 2 * sizeof(int) * nb, st>>>(&w.ctl->n_tasks[it], wp.task_cap, w.key, w.hist + YG_MAX_BUCKETS, w.perm, nb);
		else k_bin_scatter<<<sms * 4, 256, 0, st>>>(&w.ctl->n_tasks[it], wp.task_cap, w.key, w.hist + YG_MAX_BUCKETS, w.perm);
		const int *ends = w.hist + YG_MAX_BUCKETS; // the scatter kernel advanced every bucket's offset to its end
		// The per-type kernels only share atomically allocated queue slots and atomic accumulators, so they run side by side: the
		// large one fills the machine, the small ones (glass, emitters, misses) fill its tail instead of each draining the GPU alone.
		const int n_str = std::max(1, std::min(3, c->shade_streams));
		if(n_str > 1) { cudaEventRecord(c->ev_fork, st); for(int k = 0; k < n_str - 1; ++k) cudaStreamWaitEvent(c->side[k], c->ev_fork, 0); }
		int rr = 0;
		auto shadeStream = [&]() { const int k = rr++ % n_str; return k == 0 ? st : c->side[k - 1]; };
		// iteration 0 holds camera tasks only (all TASK_INTEGRATE); a DirectLight render never creates TASK_PATH tasks
		const bool has_path = it > 0 && c->scene.integ.type == YAFGPU_INT_PATHTRACING;
#define YG_SHADE_K(T, K) { k_shade<T, K><<<gs, 128, 0, shadeStream()>>>(c->scene, wp, w.pixels, in, out, w.sq, w.acc, w.ctl, it, w.perm, ends, c->n_mat_buckets, c->type_b0[T], c->type_b1[T]); ++n_launch; }
		// one kernel per (material type, task kind): +1.7 % on the Cornell bench over one kernel per type (A/B 5724 -> 5824 Mrays/s); the
		// sort key already groups by kind, so each launch covers one contiguous run
#define YG_SHADE(T) if(c->type_b1[T] >= c->type_b0[T]) { YG_SHADE_K(T, 0) if(has_path) YG_SHADE_K(T, 1) }
		YG_SHADE(YG_MAT_SD_LAMBERT) YG_SHADE(YAFGPU_MAT_SHINYDIFFUSE) YG_SHADE(YAFGPU_MAT_GLOSSY) YG_SHADE(YAFGPU_MAT_COATED_GLOSSY) YG_SHADE(YAFGPU_MAT_GLASS) YG_SHADE(YAFGPU_MAT_ROUGH_GLASS)
		YG_SHADE(YAFGPU_MAT_BLEND) YG_SHADE(YAFGPU_MAT_MIRROR) YG_SHADE(YAFGPU_MAT_NULL) YG_SHADE(YAFGPU_MAT_LIGHT)
#undef YG_SHADE
#undef YG_SHADE_K
		if(c->type_b1[YG_MAT_NODES] >= c->type_b0[YG_MAT_NODES]) { k_shade<MT_NODES, 2><<<gs, 128, 0, shadeStream()>>>(c->scene, wp, w.pixels, in, out, w.sq, w.acc, w.ctl, it, w.perm, ends, c->n_mat_buckets, c->type_b0[YG_MAT_NODES], c->type_b1[YG_MAT_NODES]); ++n_launch; }
		k_shade<MT_MISS, 2><<<gs, 128, 0, shadeStream()>>>(c->scene, wp, w.pixels, in, out, w.sq, w.acc, w.ctl, it, w.perm, ends, c->n_mat_buckets, 0, 0); ++n_launch;
		if(n_str > 1) for(int k = 0; k < n_str - 1; ++k) { cudaEventRecord(c->ev_join[k], c->side[k]); cudaStreamWaitEvent(st, c->ev_join[k], 0); }
		mark();
#define YG_SHADOW(C, R) k_shadow<C, R><<<gt, 128, 0, st>>>(c->scene, w.sq, &w.ctl->n_shadow[it], wp.shadow_cap, w.acc, &w.ctl->fetch_shadow[it], refill, c->d_counters, w.ctl)
		if(count) { if(regen_s) YG_SHADOW(true, true); else YG_SHADOW(true, false); } else { if(regen_s) YG_SHADOW(false, true); else YG_SHADOW(false, false); }
#undef YG_SHADOW
		mark();
		c->stats.kernel_launches += 4 + n_launch;
	}
	k_film_add<<<g256, 256, 0, st>>>(c->film, wp, w.pixels, w.acc, w.ctl, c->keep_nonfinite, c->d_counters, it_end);
	c->stats.kernel_launches += 1;
	YG_CUDA(c, cudaGetLastError());
	YG_CUDA(c, cudaMemcpyAsync(w.h_ctl, w.ctl, sizeof(WaveCtl), cudaMemcpyDeviceToHost, st));
	YG_CUDA(c, cudaMemcpyAsync(c->h_fault, c->d_fault, sizeof(int), cudaMemcpyDeviceToHost, st));
	YG_CUDA(c, cudaStreamSynchronize(st));
	if(int rc = faultCheck(c, false)) return rc;
	overflow = w.h_ctl->overflow != 0;
	if(!overflow && it_end < max_iters && w.h_ctl->n_tasks[it_end] != 0)
	{	// the guess was short: the film is untouched, carry on where the queue stands
		it_end = std::min(max_iters, it_end + 4);
		goto more_iterations;
	}
	c->stats.kernel_launches += 1; // k_raygen
	c->stats.waves++;
	if(!overflow)
	{	// iterations that held tasks in this wave, plus one for the next wave's guess
		int used = 0;
		for(int k = 0; k < max_iters; ++k) if(w.h_ctl->n_tasks[k] > 0) used = k + 1;
		c->iters_hint = std::min(max_iters, used + 1);
	}
	const int launched_iters = it;
	if(timed && !overflow)
	{
		for(int it = 0; it < launched_iters; ++it)
		{
			float a = 0.f, b = 0.f, d = 0.f;
			cudaEventElapsedTime(&a, ev[3 * it], ev[3 * it + 1]); cudaEventElapsedTime(&b, ev[3 * it + 1], ev[3 * it + 2]); cudaEventElapsedTime(&d, ev[3 * it + 2], ev[3 * it + 3]);
			if(w.h_ctl->n_tasks[it] > 0) { c->stats.ms_trace += a; c->stats.ms_shade += b; c->stats.launches_trace++; c->stats.launches_shade++; c->stats.rays_in_timed_trace += (uint64_t)w.h_ctl->n_tasks[it]; }
			if(w.h_ctl->n_shadow[it] > 0) { c->stats.ms_shadow += d; c->stats.launches_shadow++; c->stats.rays_in_timed_shadow += (uint64_t)w.h_ctl->n_shadow[it]; }
		}
	}
	if(overflow) { c->stats.wave_retries++; return YAFGPU_OK; }
	for(int it = 0; it < 16; ++it) { c->stats.last_wave_tasks[it] = (uint32_t)w.h_ctl->n_tasks[it]; c->stats.last_wave_shadow[it] = (uint32_t)w.h_ctl->n_shadow[it]; }
	c->stats.rays_camera += (uint64_t)wp.n_slots;
	c->stats.camera_samples += (uint64_t)wp.n_slots;
	for(int it = 1; it <= max_iters; ++it) c->stats.rays_extension += (uint64_t)w.h_ctl->n_tasks[it];
	uint64_t sh = 0;
	for(int it = 0; it < max_iters; ++it) sh += (uint64_t)w.h_ctl->n_shadow[it];
	if(c->scene.integ.tr_shad) c->stats.rays_shadow_transp += sh; else c->stats.rays_shadow += sh;
	return YAFGPU_OK;
}

int renderPass(yafgpu_ctx *c, const yafgpu_pass_desc *p)
{
	if(p->n_samples <= 0) return setErr(c, YAFGPU_ERR_INVALID, "n_samples must be positive");
	// an adaptive pass without a mask resamples every pixel: ImageFilm::doMoreSamples is true everywhere while AA_threshold is 0 (imagefilm.cc:917-920)
	int s_begin = p->sample_begin, s_end = p->sample_end;
	if(s_begin == 0 && s_end == 0) s_end = p->n_samples;
	if(s_begin < 0 || s_end > p->n_samples || s_begin >= s_end) return setErr(c, YAFGPU_ERR_INVALID, "sample range");
	size_t n_pix = 0;
	int rc = buildPixelList(c, p, n_pix);
	if(rc != YAFGPU_OK) return rc;
	if(n_pix == 0) return YAFGPU_OK;

	const IntegratorDev &ig = c->scene.integ;
	int max_add = 0;
	for(const MatDev &m : c->h_mats) max_add = std::max(max_add, m.additional_depth);
	const int rdepth = std::min(ig.r_depth + max_add, 19);
	const int max_iters = std::min(YG_MAX_ITERS, rdepth + 1 + (ig.type == YAFGPU_INT_PATHTRACING ? ig.max_bounces + 1 : 0));

	// Queue capacities per camera sample ("fans").  Tasks: the widest step of the recursion tree is the glossy x8 fan of recursiveRaytrace
	// plus the two specular children, or n_paths first path segments.  Shadow rays: one integrate() task queues, per light, ceil(samples x
	// multiplier) rays, twice that for lights that also take the BSDF-sampling half of the MIS estimator (doLightEstimation), plus the AO
	// samples; a scene with a 16-sample IBL light and a few mesh lights needs ~40 per task, and with the fixed fan of 8 used before every
	// pass first rendered and discarded whole waves (~55 % of config 4's time).
	int task_fan = 8, shadow_fan = 8;
	{
		if(ig.type == YAFGPU_INT_PATHTRACING) task_fan = std::max(task_fan, std::min(ig.n_paths + 2, 64));
		int s_int = (ig.type == YAFGPU_INT_DIRECTLIGHT && ig.do_ao) ? ig.ao_samples : 0;
		for(const LightDev &L : c->h_lights)
		{
			if(L.dirac) { s_int += 1; continue; }
			const int n = std::max(1, (int)ceilf((float)L.samples * ig.aa_light_sample_multiplier));
			s_int += n * (L.can_intersect ? 2 : 1);
		}
		shadow_fan = std::max(8, std::min(s_int + 4, 128));
	}
	if(c->task_fan_override > 0) { task_fan = c->task_fan_override; shadow_fan = std::max(shadow_fan, c->task_fan_override); }
	const size_t per_sample = (size_t)task_fan * (2 * 112) + (size_t)shadow_fan * 48 + 16;
	int wave_cap = c->wave_cap_override;
	if(wave_cap <= 0)
	{	// Default: as large as a quarter of the free HBM allows, at most 2^24 camera samples (38 GB of queues on the bench scene).  Large waves amortise the thin
		// tail iterations of a wave (a few thousand tasks per launch at depth 8+): 10^7-triangle bench frame 1 622 / 1 690 / 1 722 Mrays/s at 2^22 / 2^23 / 2^24
		// (tools/wave_sweep.sh); Cornell bench 3.57 -> 4.35 Grays/s from 2^20 to 2^23
		// (profiles/r1_wave_size.md).
		const size_t pass_samples = n_pix * (size_t)(s_end - s_begin);
		Wavefront &w0 = c->wf;
		int want = 1 << 24;
		while(want > (1 << 16) && (size_t)(want >> 1) >= pass_samples) want >>= 1; // no larger than this call needs
		if(w0.allocated && w0.wave_cap >= want && w0.task_cap == w0.wave_cap * task_fan && w0.shadow_cap == w0.wave_cap * shadow_fan) wave_cap = w0.wave_cap;   // keep what is there
		else
		{
			size_t free_b = 0, total_b = 0;
			cudaMemGetInfo(&free_b, &total_b);
			if(w0.allocated) free_b += (size_t)w0.task_cap * (2 * 112) + (size_t)w0.shadow_cap * 48 + (size_t)w0.wave_cap * 16;
			wave_cap = want;
			while(wave_cap > (1 << 16) && (size_t)wave_cap * per_sample > free_b / 4) wave_cap >>= 1;
		}
	}
	cudaError_t e = allocWavefront(c, wave_cap, wave_cap * task_fan, wave_cap * shadow_fan);
	if(e != cudaSuccess) return cudaErr(c, e, "wavefront allocation");
	Wavefront &w = c->wf;

	cudaEventRecord(c->ev0, c->stream);
	const int n_smp = s_end - s_begin;
	size_t pix0 = 0;
	int smp0 = 0;
	// a wave size that overflowed once on this scene is not tried again (the scene decides the fan-out, not the pass)
	int cur_cap = (c->wave_cur_cap > 0 && c->wave_cur_cap < wave_cap) ? c->wave_cur_cap : wave_cap;
	while(smp0 < n_smp)
	{
		WaveParams wp;
		// fill a wave: whole-pixel-list x several samples when the list is small, else a slice of the list for one sample
		if((size_t)cur_cap >= n_pix && pix0 == 0)
		{
			wp.n_pix = (int)n_pix; wp.pix_base = 0;
			const int ns = std::max(1, std::min(n_smp - smp0, (int)((size_t)cur_cap / n_pix)));
			wp.n_slots = wp.n_pix * ns;
		}
		else
		{
			wp.pix_base = (int)pix0; wp.n_pix = (int)std::min((size_t)cur_cap, n_pix - pix0);
			wp.n_slots = wp.n_pix;
		}
		wp.sample_base = s_begin + smp0;
		wp.n_samples_pass = p->n_samples; wp.pass_offs = p->offset; wp.aa_passes = c->film.aa_passes;
		wp.task_cap = w.task_cap; wp.shadow_cap = w.shadow_cap;
		wp.seed = hashSeed((unsigned)p->offset, 0x2545F491u, (unsigned)p->aa_pass_number);   // per pass, NOT per wave: a sample's random numbers must not depend on how the pass is cut into waves
		bool overflow = false;
		rc = runWave(c, wp, max_iters, overflow);
		if(rc != YAFGPU_OK) return rc;
		if(overflow)
		{
			if(cur_cap <= 1024) return setErr(c, YAFGPU_ERR_NOMEM, "wavefront queues overflow even at the smallest wave size");
			cur_cap /= 2;
			c->wave_cur_cap = cur_cap;
			continue; // retry the same range with a smaller wave
		}
		const int ns_done = wp.n_slots / wp.n_pix;
		if(wp.pix_base == 0 && (size_t)wp.n_pix == n_pix) smp0 += ns_done;
		else
		{
			pix0 += (size_t)wp.n_pix;
			if(pix0 >= n_pix) { pix0 = 0; smp0 += 1; }
		}
	}
	cudaEventRecord(c->ev1, c->stream);
	cudaStreamSynchronize(c->stream);
	float ms = 0.f; cudaEventElapsedTime(&ms, c->ev0, c->ev1);
	c->stats.ms_total += ms;
	return YAFGPU_OK;
}

} // namespace yg
