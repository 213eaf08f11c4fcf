// K3/K4 device traversal: closest hit and any hit over the BVH2 of scene_dev.cuh.
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
#pragma once
#include "scene_dev.cuh"

namespace yg {

struct RayD { V3 o; float tmin; V3 d; float tmax; };
struct HitD { float t, u, v; int tri; };

#define YG_STACK 96

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

// slab test of one child box against [tlo, thi]; conservative (NaN-safe min/max, far side widened by 2 ulp-ish)
YG_D bool slab(float bx0, float by0, float bz0, float bx1, float by1, float bz1, const V3 o, const V3 id, float tlo, float thi, float &tnear)
{
	float t0 = (bx0 - o.x) * id.x, t1 = (bx1 - o.x) * id.x;
	float tn = fminf(t0, t1), tf = fmaxf(t0, t1);
	t0 = (by0 - o.y) * id.y; t1 = (by1 - o.y) * id.y;
	tn = fmaxf(tn, fminf(t0, t1)); tf = fminf(tf, fmaxf(t0, t1));
	t0 = (bz0 - o.z) * id.z; t1 = (bz1 - o.z) * id.z;
	tn = fmaxf(tn, fminf(t0, t1)); tf = fminf(tf, fmaxf(t0, t1));
	tf *= 1.0000004f;
	tnear = tn;
	return tn <= tf && tf >= tlo && tn <= thi;
}

struct TravCounters { unsigned long long nodes, tris; };

// A traversal stack that is full cannot take the far child: the subtree would be skipped and the image silently wrong.  The entry is
// still dropped (no memory is touched out of bounds) but the accelerator's fault word is raised; every host entry point that
// synchronises (yafgpu_trace, yafgpu_render_pass, yafgpu_sync) reads it and fails with YAFGPU_ERR_CUDA instead of returning the result.
YG_D void stackFault(const Accel &acc) { if(acc.fault) *acc.fault = 1; }

// Closest hit. `dist` = ray.tmax < 0 ? inf : ray.tmax (scene.cc:900-901).
template<bool COUNT>
YG_D HitD traceClosest(const Accel &acc, const RayD &r, TravCounters *cnt)
{
	HitD h; h.t = 0.f; h.u = 0.f; h.v = 0.f; h.tri = -1;
	float z = (r.tmax < 0.f) ? __int_as_float(0x7f800000) : r.tmax;
	const V3 id = mk(1.f / r.d.x, 1.f / r.d.y, 1.f / r.d.z);
	const float tlo = r.tmin * 0.99f;
	int stack[YG_STACK];
	int sp = 0;
	int cur = acc.root;
	if(acc.n_tris == 0) return h;
	for(;;)
	{
		if(cur >= 0)
		{
			const BvhNode *nd = acc.nodes + cur;
			const float4 n0 = __ldg(&nd->n0), n1 = __ldg(&nd->n1), n2 = __ldg(&nd->n2), n3 = __ldg(&nd->n3);
			if(COUNT) cnt->nodes++;
			float ta, tb;
			const bool ha = slab(n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, r.o, id, tlo, z, ta);
			const bool hb = slab(n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, r.o, id, tlo, z, tb);
			const int ca = __float_as_int(n3.x), cb = __float_as_int(n3.y);
			if(ha && hb)
			{
				const bool a_first = ta <= tb;
				cur = a_first ? ca : cb;
				if(sp < acc.stack_limit) stack[sp++] = a_first ? cb : ca; else stackFault(acc);
				continue;
			}
			if(ha) { cur = ca; continue; }
			if(hb) { cur = cb; continue; }
		}
		else
		{
			const int code = ~cur;
			const int first = code >> 3, count = (code & 7) + 1;
			for(int k = 0; k < count; ++k)
			{
				const TriAccel *tp = acc.tris + first + k;
				const float4 v0e = __ldg(&tp->v0e), e1i = __ldg(&tp->e1i), e2f = __ldg(&tp->e2f);
				if(COUNT) cnt->tris++;
				float t, u, v;
				if(!triIntersect(v0e, e1i, e2f, r.o, r.d, t, u, v)) continue;
				if(!(__float_as_int(e2f.w) & TRI_VIS_CAMERA)) continue;
				const int id_ = __float_as_int(e1i.w);
				if(t >= r.tmin && (t < z || (t == z && h.tri >= 0 && id_ < h.tri)))
				{
					z = t; h.t = t; h.u = u; h.v = v; h.tri = id_;
				}
			}
		}
		if(sp == 0) break;
		cur = stack[--sp];
	}
	return h;
}

// Any hit for opaque shadows. The caller passes the ALREADY SHIFTED origin and dist (scene.cc:965-970).
// Returns the global id of a blocker (any), or -1.
template<bool COUNT>
YG_D int traceShadow(const Accel &acc, const V3 o, const V3 d, float dist, TravCounters *cnt)
{
	if(acc.n_tris == 0) return -1;
	const V3 id = mk(1.f / d.x, 1.f / d.y, 1.f / d.z);
	int stack[YG_STACK];
	int sp = 0;
	int cur = acc.root;
	for(;;)
	{
		if(cur >= 0)
		{
			const BvhNode *nd = acc.nodes + cur;
			const float4 n0 = __ldg(&nd->n0), n1 = __ldg(&nd->n1), n2 = __ldg(&nd->n2), n3 = __ldg(&nd->n3);
			if(COUNT) cnt->nodes++;
			float ta, tb;
			const bool ha = slab(n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, o, id, 0.f, dist, ta);
			const bool hb = slab(n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, o, id, 0.f, dist, tb);
			const int ca = __float_as_int(n3.x), cb = __float_as_int(n3.y);
			if(ha && hb) { cur = ca; if(sp < acc.stack_limit) stack[sp++] = cb; else stackFault(acc); continue; }
			if(ha) { cur = ca; continue; }
			if(hb) { cur = cb; continue; }
		}
		else
		{
			const int code = ~cur;
			const int first = code >> 3, count = (code & 7) + 1;
			for(int k = 0; k < count; ++k)
			{
				const TriAccel *tp = acc.tris + first + k;
				const float4 v0e = __ldg(&tp->v0e), e1i = __ldg(&tp->e1i), e2f = __ldg(&tp->e2f);
				if(COUNT) cnt->tris++;
				float t, u, v;
				if(!triIntersect(v0e, e1i, e2f, o, d, t, u, v)) continue;
				if(t < dist && t >= 0.f && (__float_as_int(e2f.w) & TRI_VIS_SHADOW)) return __float_as_int(e1i.w);
			}
		}
		if(sp == 0) break;
		cur = stack[--sp];
	}
	return -1;
}

// ---- warp-persistent traversal with ray regeneration ------------------------------------------------------------------------
// The wavefront kernels do not map ray i to thread i.  Every warp keeps 32 rays in flight; a lane whose ray has finished stores
// its result and takes the next unprocessed ray from a global counter (one warp-aggregated atomicAdd per refill), so lanes do not
// idle until the slowest ray of their warp is done.  Inside, traversal is "while-while": all lanes first descend through inner
// nodes until each holds a leaf (or has nothing left), then all process their leaf together, so the node loop and the
// Möller–Trumbore loop each run with most lanes active instead of interleaving per lane.
//
// IO supplies:  void load(int idx, V3 &o, V3 &d, float &tmin, float &dist)
//               void store(int idx, const HitD &h, unsigned lanes_storing)   called warp-synchronously by the lanes in `lanes_storing`
// ANYHIT: Scene::isShadowed semantics (origin already shifted by load, accept t in [0, dist), TRI_VIS_SHADOW, first hit ends the ray);
// else Scene::intersect semantics (accept t in [tmin, z), TRI_VIS_CAMERA, lowest id at equal t).
#define YG_EMPTY ((int)0x80000000)
#ifndef YG_REGEN_THRESH
#define YG_REGEN_THRESH 22 /* refill as soon as fewer lanes than this are still traversing (tuned on random rays in the 10^7-triangle mesh) */
#endif
#ifndef YG_REGEN_ROUNDS
#define YG_REGEN_ROUNDS 6
#endif

template<bool ANYHIT, bool COUNT, class IO>
__device__ __forceinline__ void traversePersistent(const Accel &acc, int n, int *__restrict__ next, IO &io, TravCounters *cnt, const int refill_below = YG_REGEN_THRESH)
{
	const unsigned FULL = 0xffffffffu;
	const int lane = threadIdx.x & 31;
	const unsigned lt_mask = (1u << lane) - 1u;
	int ray = -1;          // index of the ray this lane works on, -1 = none
	int cur = YG_EMPTY;    // node to visit next: >= 0 inner node, < 0 leaf code, YG_EMPTY = traversal finished
	V3 o = mk(0.f), d = mk(0.f), id = mk(0.f);
	float tmin = 0.f, z = 0.f, tlo = 0.f;
	HitD h; h.t = 0.f; h.u = 0.f; h.v = 0.f; h.tri = -1;
	int stack[YG_STACK];
	int sp = 0;
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
					float dist;
					io.load(idx, o, d, tmin, dist);
					ray = idx;
					id = mk(1.f / d.x, 1.f / d.y, 1.f / d.z);
					z = dist; tlo = ANYHIT ? 0.f : tmin * 0.99f;
					h.t = 0.f; h.u = 0.f; h.v = 0.f; h.tri = -1;
					sp = 0; cur = acc.n_tris != 0 ? acc.root : YG_EMPTY;
				}
			}
		}
		if(__all_sync(FULL, ray < 0)) { if(exhausted) break; else continue; }

		// ---- a few while-while rounds, then look for lanes to refill ----
		#pragma unroll 1
		for(int round = 0; round < YG_REGEN_ROUNDS; ++round)
		{
			// node phase: descend until this lane holds a leaf or has finished
			while(cur >= 0)
			{
				const BvhNode *nd = acc.nodes + cur;
				const float4 n0 = __ldg(&nd->n0), n1 = __ldg(&nd->n1), n2 = __ldg(&nd->n2), n3 = __ldg(&nd->n3);
				if(COUNT) cnt->nodes++;
				float ta, tb;
				const bool ha = slab(n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, o, id, tlo, z, ta);
				const bool hb = slab(n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, o, id, tlo, z, tb);
				const int ca = __float_as_int(n3.x), cb = __float_as_int(n3.y);
				if(ha && hb)
				{
					const bool a_first = ANYHIT ? true : (ta <= tb);
					cur = a_first ? ca : cb;
					if(sp < acc.stack_limit) stack[sp++] = a_first ? cb : ca; else stackFault(acc);
				}
				else if(ha) cur = ca;
				else if(hb) cur = cb;
				else cur = sp > 0 ? stack[--sp] : YG_EMPTY;
			}
			// leaf phase
			if(cur != YG_EMPTY)
			{
				const int code = ~cur;
				const int first = code >> 3, count = (code & 7) + 1;
				bool ended = false;
				for(int k = 0; k < count; ++k)
				{
					const TriAccel *tp = acc.tris + first + k;
					const float4 v0e = __ldg(&tp->v0e), e1i = __ldg(&tp->e1i), e2f = __ldg(&tp->e2f);
					if(COUNT) cnt->tris++;
					float t, u, v;
					if(!triIntersect(v0e, e1i, e2f, o, d, t, u, v)) continue;
					const int fl = __float_as_int(e2f.w), id_ = __float_as_int(e1i.w);
					if(ANYHIT)
					{
						if(t < z && t >= 0.f && (fl & TRI_VIS_SHADOW)) { h.tri = id_; h.t = 1.f; ended = true; break; }
					}
					else
					{
						if(!(fl & TRI_VIS_CAMERA)) continue;
						if(t >= tmin && (t < z || (t == z && h.tri >= 0 && id_ < h.tri))) { z = t; h.t = t; h.u = u; h.v = v; h.tri = id_; }
					}
				}
				cur = (ended || sp == 0) ? YG_EMPTY : stack[--sp];
			}
			// leave early when enough lanes have nothing to do
			if(__popc(__ballot_sync(FULL, ray >= 0 && cur != YG_EMPTY)) < refill_below) break;
		}

		// ---- store finished rays (warp-synchronous) ----
		const bool fin = ray >= 0 && cur == YG_EMPTY;
		const unsigned fin_mask = __ballot_sync(FULL, fin);
		if(fin) { io.store(ray, h, fin_mask); ray = -1; }
	}
}

} // namespace yg
