// K3/K4 device traversal: closest hit and any hit over the W-wide BVH of scene_dev.cuh.
//
// Semantics follow the reference, not a generic ray tracer:
//   closest : Scene::intersect -> TriKdTree::intersect (/root/reference/src/common/scene.cc:896-927, kdtree_triangle.cc:684-837)
//             accept t_hit < z && t_hit >= ray.tmin && visibility in {Normal, VisibleNoShadows}; z starts at dist.
//   shadow  : Scene::isShadowed -> TriKdTree::intersectS (scene.cc:962-995, kdtree_triangle.cc:840-977)
//             caller has already moved the origin by dir*tmin and set dist = tmax - 2*tmin; accept t_hit < dist && t_hit >= 0
//             && visibility in {Normal, InvisibleShadowsOnly}.
//   triangle: Triangle::intersect (/root/reference/include/common/triangle.h:223-259), float32 Möller–Trumbore on the cached
//             a, edge_1, edge_2, eps, same operation order, no FMA contraction (-fmad=false).
// Equal-t candidates: the kd-tree keeps the first one it visits (strict <); a BVH visits in another order, so the tie is
// broken deterministically towards the LOWER global triangle id (documented tie rule).
//
// Only the triangle test is parity-critical; the box tests merely cull and are free to be fast: the near / far plane of every child
// is fetched by address from the direction signs and tested in t-space with one fused multiply-add (t = plane / d - o / d), children
// are visited nearest first, the rest pushed farthest first with their entry distance so that a popped entry beyond the current hit
// costs no node fetch.  The loop is "while-while" (Aila & Laine 2009): all lanes first descend through inner nodes until each holds
// a leaf, then run Möller–Trumbore together — the kernel is bound by instruction issue at ~10 of 32 active lanes (profiles/r2b_*),
// so keeping the two code paths apart is worth more than anything that saves bytes.
// The first YG_STACK_SHARED entries of the per-ray stack live in shared memory (one 8-byte column per thread), the tail in local
// memory; a full stack raises Accel::fault (the host turns it into an error) instead of dropping the subtree silently.
#pragma once
#include "scene_dev.cuh"

namespace yg {

struct RayD { V3 o; float tmin; V3 d; float tmax; };
struct HitD { float t, u, v; int tri; };

#ifndef YG_BVH_WIDTH
#define YG_BVH_WIDTH 4
#endif
#define YG_STACK 64          /* entries per ray: link + entry distance, 8 B each */
#ifndef YG_STACK_SHARED
#define YG_STACK_SHARED 16   /* of which in shared memory in the kernels that provide it (16 x 8 B x 128 threads = 16 KB per block) */
#endif
#define YG_NODE_U4 (YG_BVH_WIDTH == 4 ? 8 : 16)

// Möller–Trumbore exactly as triangle.h:223-259. Returns true and t,u,v when the reference would return true.
YG_D bool triIntersect(const float4 v0e, const float4 e1i, const float4 e2f, const V3 o, const V3 d, float &t, float &u, float &v)
{
	const V3 e1 = mk(e1i.x, e1i.y, e1i.z), e2 = mk(e2f.x, e2f.y, e2f.z);
	const V3 pvec = cross(d, e2);
	const float det = dot(e1, pvec);
	const float eps = v0e.w;
	if(det > -eps && det < eps) return false;
	const float inv_det = 1.f / det;
	const V3 tvec = mk(o.x - v0e.x, o.y - v0e.y, o.z - v0e.z);
	u = dot(tvec, pvec) * inv_det;
	if(u < 0.f || u > 1.f) return false;
	const V3 qvec = cross(tvec, e1);
	v = dot(d, qvec) * inv_det;
	if((v < 0.f) || ((u + v) > 1.f)) return false;
	t = dot(e2, qvec) * inv_det;
	if(t < eps) return false;
	return true;
}

struct TravCounters { unsigned long long nodes, tris; };

// A traversal stack that is full cannot take another child: the subtree would be skipped and the image silently wrong.  The entry is
// still dropped (no memory is touched out of bounds) but the accelerator's fault word is raised; every host entry point that
// synchronises (yafgpu_trace, yafgpu_render_pass, yafgpu_sync) reads it and fails with YAFGPU_ERR_CUDA instead of returning the result.
YG_D void stackFault(const Accel &acc) { if(acc.fault) *acc.fault = 1; }

// Per-ray stack: entries [0, NS) in shared memory (column of this thread: s[i * stride]), the rest in local memory.  NS = 0: all local.
template<int NS>
struct TravStack
{
	uint2 *s; int stride;
	uint2 l[YG_STACK - NS];
	__device__ __forceinline__ void put(int i, uint2 v) { if(NS > 0 && i < NS) s[i * stride] = v; else l[i - NS] = v; }
	__device__ __forceinline__ uint2 get(int i) const { return (NS > 0 && i < NS) ? s[i * stride] : l[i - NS]; }
};

// 1/d for the box tests only: a zero component becomes +-1e-30 so that no inf/NaN enters the fused arithmetic (the box of a ray
// parallel to an axis is then tested with slopes of 1e30, which decides "inside the slab or not" exactly as it should)
__device__ __forceinline__ float safeInv(float d) { return 1.f / (fabsf(d) > 1e-30f ? d : copysignf(1e-30f, d)); }

// one 32-byte chunk through the read-only path: LDG.E.256 on sm_100a
struct __align__(32) F8 { float v[8]; };
__device__ __forceinline__ F8 ld256(const void *p)
{
	F8 r;
	asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
	             : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7]) : "l"(p));
	return r;
}

// Per-ray traversal state: the constants of the box test, the search interval and the stack position.
struct TravRay
{
	V3 o, d;
	float idx, idy, idz, ox, oy, oz;   // 1/d (safe) and -o/d
	bool px, py, pz;                   // direction signs: the near plane of an axis is its lo plane when positive
	float tlo, z;                      // box tests clip to [tlo, z]; z is the far limit of the search (the closest hit so far)
	int cur, sp;
	__device__ __forceinline__ void start(const Accel &acc, const V3 o_, const V3 d_, float tlo_, float z_)
	{
		o = o_; d = d_;
		idx = safeInv(d.x); idy = safeInv(d.y); idz = safeInv(d.z);
		ox = -o.x * idx; oy = -o.y * idy; oz = -o.z * idz;
		px = idx >= 0.f; py = idy >= 0.f; pz = idz >= 0.f;
		tlo = fmaxf(tlo_, 0.f); z = z_;
		sp = 0; cur = acc.n_tris != 0 ? acc.root : YG_NO_CHILD;
	}
};

// nearest pending subtree that can still hold something closer than z, or YG_NO_CHILD
template<int NS>
__device__ __forceinline__ int popNext(TravRay &r, const TravStack<NS> &st)
{
	while(r.sp > 0)
	{
		const uint2 e = st.get(--r.sp);
		if(__uint_as_float(e.y) <= r.z * 1.0000004f) return (int)e.x;
	}
	return YG_NO_CHILD;
}

// One inner node (r.cur >= 0): box tests of its W children; r.cur becomes the nearest hit child (ORDERED) or the first one, the others
// are pushed (farthest first when ORDERED); with no child hit the next pending subtree is popped.
// The ordering is branch-free: (entry distance | slot, link) pairs go through a sorting network, misses sort last, then W - 1
// predicated pushes.  (A loop "extract the farthest, push it" ran at 3.3 of 32 lanes and took 17 % of the kernel's issue slots,
// profiles/r2c_ncu_k_trace_bvh4_fp32_regen.txt.)
#define YG_CX(a, b) { const bool sw_ = key[b] < key[a]; const unsigned ka_ = sw_ ? key[b] : key[a], kb_ = sw_ ? key[a] : key[b]; \
                      const int la_ = sw_ ? link[b] : link[a], lb_ = sw_ ? link[a] : link[b]; key[a] = ka_; key[b] = kb_; link[a] = la_; link[b] = lb_; }
template<bool ORDERED, bool COUNT, int NS>
__device__ __forceinline__ void nodeStep(const Accel &acc, TravRay &r, TravStack<NS> &st, TravCounters *cnt)
{
	constexpr int W = YG_BVH_WIDTH, V = W / 4;
	const char *nd = (const char *)(acc.nodes + (size_t)r.cur * YG_NODE_U4);
	if(COUNT) cnt->nodes++;
	const float zf = r.z * 1.0000004f;
	const int limit = acc.stack_limit;
	unsigned key[W];   // entry distance bits with the slot in the low 3 bits; 0xffffffff = missed
	int link[W];
#pragma unroll
	for(int v = 0; v < V; ++v)
	{
		const F8 cx = ld256(nd + 32 * v), cy = ld256(nd + 32 * (V + v)), cz = ld256(nd + 32 * (2 * V + v)), cl = ld256(nd + 32 * (3 * V + v));
#pragma unroll
		for(int j = 0; j < 4; ++j)
		{
			const int k = 4 * v + j;
			// near / far plane of each axis by the direction sign
			const float nxp = r.px ? cx.v[j] : cx.v[4 + j], fxp = r.px ? cx.v[4 + j] : cx.v[j];
			const float nyp = r.py ? cy.v[j] : cy.v[4 + j], fyp = r.py ? cy.v[4 + j] : cy.v[j];
			const float nzp = r.pz ? cz.v[j] : cz.v[4 + j], fzp = r.pz ? cz.v[4 + j] : cz.v[j];
			const float tn = fmaxf(fmaxf(fmaf(nxp, r.idx, r.ox), fmaf(nyp, r.idy, r.oy)), fmaxf(fmaf(nzp, r.idz, r.oz), r.tlo));
			const float tf = fminf(fminf(fmaf(fxp, r.idx, r.ox), fmaf(fyp, r.idy, r.oy)), fminf(fmaf(fzp, r.idz, r.oz), zf));
			link[k] = __float_as_int(cl.v[j]);
			// the boxes carry an absolute pad of 2^-19 of the scene size (bvh_build.cu), which covers the rounding of these six products;
			// an absent child has an inverted box and fails here
			key[k] = (tn <= tf) ? ((__float_as_uint(tn) & ~7u) | (unsigned)k) : 0xffffffffu;
		}
	}
	if(ORDERED)
	{
		if(W == 4) { YG_CX(0, 1) YG_CX(2, 3) YG_CX(0, 2) YG_CX(1, 3) YG_CX(1, 2) }
		else
		{	// 19-comparator network for 8 keys
			YG_CX(0, 1) YG_CX(2, 3) YG_CX(W - 4, W - 3) YG_CX(W - 2, W - 1) YG_CX(0, 2) YG_CX(1, 3) YG_CX(W - 4, W - 2) YG_CX(W - 3, W - 1) YG_CX(1, 2) YG_CX(W - 3, W - 2)
			YG_CX(0, W - 4) YG_CX(1, W - 3) YG_CX(2, W - 2) YG_CX(3, W - 1) YG_CX(2, W - 4) YG_CX(3, W - 3) YG_CX(1, 2) YG_CX(3, W - 4) YG_CX(W - 3, W - 2)
		}
#pragma unroll
		for(int k = W - 1; k >= 1; --k)
			if(key[k] != 0xffffffffu) { if(r.sp < limit) st.put(r.sp++, make_uint2((unsigned)link[k], key[k] & ~7u)); else stackFault(acc); }
		r.cur = key[0] != 0xffffffffu ? link[0] : popNext(r, st);
	}
	else
	{
		int next = YG_NO_CHILD;
#pragma unroll
		for(int k = 0; k < W; ++k)
		{
			if(key[k] == 0xffffffffu) continue;
			if(next == YG_NO_CHILD) next = link[k];
			else if(r.sp < limit) st.put(r.sp++, make_uint2((unsigned)link[k], key[k] & ~7u)); else stackFault(acc);
		}
		r.cur = next != YG_NO_CHILD ? next : popNext(r, st);
	}
}
#undef YG_CX

// One leaf link (< 0, not YG_NO_CHILD): Triangle::intersect on its triangles; f decides.  Returns true when f ended the search.
template<bool COUNT, class F>
__device__ __forceinline__ bool leafTest(const Accel &acc, TravRay &r, int link, F &f, TravCounters *cnt)
{
	const int code = ~link;
	const int first = code >> 3, count = (code & 7) + 1;
	for(int k = 0; k < count; ++k)
	{
		const TriAccel *tp = acc.tris + first + k;
		const float4 v0e = __ldg(&tp->v0e), e1i = __ldg(&tp->e1i), e2f = __ldg(&tp->e2f);
		if(COUNT) cnt->tris++;
		float t, u, v;
		if(!triIntersect(v0e, e1i, e2f, r.o, r.d, t, u, v)) continue;
		if(f(e1i, e2f, t, u, v, r.z)) return true;
	}
	return false;
}

// Node phase of one lane: descend until the lane is blocked on a leaf.  With POSTPONE the first leaf met is set aside (`stash`) and the
// descent goes on with the next pending subtree, so that the lane stays in the node loop with the rest of its warp instead of idling
// until every other lane has found a leaf of its own (the node loop ran at 11.5 of 32 lanes without this, profiles/r2c_*); the leaf
// test is order-independent (the closest hit is the minimum over all accepted triangles), only the shrinking of z is delayed.
#ifndef YG_POSTPONE
#define YG_POSTPONE 1
#endif
template<bool ORDERED, bool COUNT, bool POSTPONE, int NS>
__device__ __forceinline__ void nodePhase(const Accel &acc, TravRay &r, int &stash, TravStack<NS> &st, TravCounters *cnt)
{
	for(;;)
	{
		if(r.cur >= 0) nodeStep<ORDERED, COUNT>(acc, r, st, cnt);
		else if(POSTPONE && r.cur != YG_NO_CHILD && stash == YG_NO_CHILD) { stash = r.cur; r.cur = popNext(r, st); }
		else break;
	}
}

// Leaf phase of one lane: the postponed leaf, then the one in hand.  Returns true when f ended the search.
template<bool COUNT, int NS, class F>
__device__ __forceinline__ bool leafPhase(const Accel &acc, TravRay &r, int &stash, TravStack<NS> &st, F &f, TravCounters *cnt)
{
	if(stash != YG_NO_CHILD)
	{
		const int l = stash; stash = YG_NO_CHILD;
		if(leafTest<COUNT>(acc, r, l, f, cnt)) { r.cur = YG_NO_CHILD; return true; }
	}
	if(r.cur != YG_NO_CHILD && r.cur < 0)
	{
		if(leafTest<COUNT>(acc, r, r.cur, f, cnt)) { r.cur = YG_NO_CHILD; return true; }
		r.cur = popNext(r, st);
	}
	return false;
}

// One ray, start to finish ("while-while").  F: bool f(e1i, e2f, t, u, v, float &z) — called for every triangle Triangle::intersect
// accepts; may lower z; returns true to end the traversal.  ORDERED: visit children nearest first (closest hit); else in slot order.
template<bool ORDERED, bool COUNT, int NS, class F>
__device__ __forceinline__ void traverseWide(const Accel &acc, const V3 o, const V3 d, float tlo, float &z, TravStack<NS> &st, F &f, TravCounters *cnt)
{
	TravRay r;
	r.start(acc, o, d, tlo, z);
	int stash = YG_NO_CHILD;
	while(r.cur != YG_NO_CHILD)
	{
		nodePhase<ORDERED, COUNT, ORDERED && YG_POSTPONE>(acc, r, stash, st, cnt);
		if(leafPhase<COUNT>(acc, r, stash, st, f, cnt)) break;
	}
	z = r.z;
}

// ---- the two searches of the reference --------------------------------------------------------------------------------------------
struct ClosestF
{
	float tmin; HitD h;
	__device__ __forceinline__ bool operator()(const float4 e1i, const float4 e2f, float t, float u, float v, float &z)
	{
		if(!(__float_as_int(e2f.w) & TRI_VIS_CAMERA)) return false;
		const int id_ = __float_as_int(e1i.w);
		if(t >= tmin && (t < z || (t == z && h.tri >= 0 && id_ < h.tri))) { z = t; h.t = t; h.u = u; h.v = v; h.tri = id_; }
		return false;
	}
};

struct AnyF
{
	int tri;
	__device__ __forceinline__ bool operator()(const float4 e1i, const float4 e2f, float t, float, float, float &z)
	{
		if(t < z && t >= 0.f && (__float_as_int(e2f.w) & TRI_VIS_SHADOW)) { tri = __float_as_int(e1i.w); return true; }
		return false;
	}
};

// Closest hit. `dist` = ray.tmax < 0 ? inf : ray.tmax (scene.cc:900-901).
template<bool COUNT, int NS>
__device__ __forceinline__ HitD traceClosestS(const Accel &acc, const RayD &r, TravStack<NS> &st, TravCounters *cnt)
{
	ClosestF f; f.tmin = r.tmin; f.h.t = 0.f; f.h.u = 0.f; f.h.v = 0.f; f.h.tri = -1;
	float z = (r.tmax < 0.f) ? __int_as_float(0x7f800000) : r.tmax;
	traverseWide<true, COUNT, NS>(acc, r.o, r.d, r.tmin * 0.99f, z, st, f, cnt);
	return f.h;
}

// Any hit for opaque shadows. The caller passes the ALREADY SHIFTED origin and dist (scene.cc:965-970).
// Returns the global id of a blocker (any), or -1.
template<bool COUNT, int NS>
__device__ __forceinline__ int traceShadowS(const Accel &acc, const V3 o, const V3 d, float dist, TravStack<NS> &st, TravCounters *cnt)
{
	AnyF f; f.tri = -1;
	float z = dist;
	traverseWide<false, COUNT, NS>(acc, o, d, 0.f, z, st, f, cnt);
	return f.tri;
}

// all-local-stack flavours for code that runs outside the traversal kernels (mesh-light intersection inside the shading kernels)
template<bool COUNT>
YG_D HitD traceClosest(const Accel &acc, const RayD &r, TravCounters *cnt)
{
	TravStack<0> st; st.s = nullptr; st.stride = 0;
	return traceClosestS<COUNT, 0>(acc, r, st, cnt);
}

// ---- warp-persistent traversal with ray regeneration ------------------------------------------------------------------------
// The kernel is bound by instruction issue, and a warp that maps ray i to lane i issues for as long as its SLOWEST ray lives (node
// visits per ray spread from a handful to ~100).  Here every warp keeps 32 rays in flight instead: a lane whose ray has finished
// stores its result and takes the next unprocessed ray from a global counter (one warp-aggregated atomicAdd per refill), as soon as
// fewer than `refill_below` lanes are still busy.  Inside, the loop is the same while-while as traverseWide.
//
// IO supplies:  void load(int idx, V3 &o, V3 &d, float &tmin, float &dist)
//               void store(int idx, const HitD &h, unsigned lanes_storing)   called warp-synchronously by the lanes in `lanes_storing`
// ANYHIT: Scene::isShadowed semantics (origin already shifted by load, accept t in [0, dist), TRI_VIS_SHADOW, first hit ends the ray;
// a blocked ray reports h.tri >= 0); else Scene::intersect semantics.
#ifndef YG_REGEN_ROUNDS
#define YG_REGEN_ROUNDS 4
#endif
template<bool ANYHIT, bool COUNT, int NS, class IO>
__device__ __forceinline__ void traversePersistent(const Accel &acc, int n, int *__restrict__ next, IO &io, TravStack<NS> &st, TravCounters *cnt, const int refill_below)
{
	const unsigned FULL = 0xffffffffu;
	const int lane = threadIdx.x & 31;
	const unsigned lt_mask = (1u << lane) - 1u;
	int ray = -1;          // index of the ray this lane works on, -1 = none
	int stash = YG_NO_CHILD;
	TravRay r; r.cur = YG_NO_CHILD; r.sp = 0;
	ClosestF fc; fc.tmin = 0.f; fc.h.t = 0.f; fc.h.u = 0.f; fc.h.v = 0.f; fc.h.tri = -1;
	AnyF fa; fa.tri = -1;
	bool exhausted = false; // the global counter has run past n
	for(;;)
	{
		// ---- refill idle lanes (warp-synchronous) ----
		const unsigned idle = __ballot_sync(FULL, ray < 0);
		if(idle && !exhausted)
		{
			const int want = __popc(idle);
			int base = 0;
			if(lane == 0) base = atomicAdd(next, want);
			base = __shfl_sync(FULL, base, 0);
			if(base + want >= n) exhausted = true;
			if(ray < 0)
			{
				const int idx = base + __popc(idle & lt_mask);
				if(idx < n)
				{
					V3 o, d; float tmin, dist;
					io.load(idx, o, d, tmin, dist);
					ray = idx;
					r.start(acc, o, d, ANYHIT ? 0.f : tmin * 0.99f, dist);
					fc.tmin = tmin; fc.h.t = 0.f; fc.h.u = 0.f; fc.h.v = 0.f; fc.h.tri = -1; fa.tri = -1;
				}
			}
		}
		if(__all_sync(FULL, ray < 0)) { if(exhausted) break; else continue; }

		// ---- a few while-while rounds, then look for lanes to refill ----
#pragma unroll 1
		for(int round = 0; round < YG_REGEN_ROUNDS; ++round)
		{
			nodePhase<!ANYHIT, COUNT, !ANYHIT && YG_POSTPONE>(acc, r, stash, st, cnt);
			if(ANYHIT) leafPhase<COUNT>(acc, r, stash, st, fa, cnt); else leafPhase<COUNT>(acc, r, stash, st, fc, cnt);
			if(__popc(__ballot_sync(FULL, ray >= 0 && r.cur != YG_NO_CHILD)) < refill_below) break;
		}

		// ---- store finished rays (warp-synchronous) ----
		const bool fin = ray >= 0 && r.cur == YG_NO_CHILD;
		const unsigned fin_mask = __ballot_sync(FULL, fin);
		if(fin)
		{
			HitD h = fc.h;
			if(ANYHIT) { h.tri = fa.tri; h.t = fa.tri >= 0 ? 1.f : 0.f; }
			io.store(ray, h, fin_mask);
			ray = -1;
		}
	}
}

// this thread's column of a block-wide shared stack area of YG_STACK_SHARED * blockDim.x entries
__device__ __forceinline__ void bindSharedStack(TravStack<YG_STACK_SHARED> &st, uint2 *area) { st.s = area + threadIdx.x; st.stride = blockDim.x; }

} // namespace yg
