// K1 — GPU BVH build (replaces TriKdTree::TriKdTree + buildTree, /root/reference/src/common/kdtree_triangle.cc:76-157,468-675).
//
// Only the CLOSEST HIT has to agree with the reference, not the tree, so the accelerator is a linear BVH:
//   tri_setup   : per triangle a, edge_1, edge_2, eps exactly as Triangle::updateIntersectionCachedValues
//                 (triangle.h:197-207), geometric normal as Triangle::recNormal (triangle.h:274-281), AABB, centroid
//   scene bound : atomic min/max reduction (float ordered-int trick)
//   morton      : 63-bit codes (21 bits/axis) of the centroids
//   sort        : cub::DeviceRadixSort (library primitive; everything else here is hand-written)
//   hierarchy   : Karras 2012 "Maximizing parallelism in the construction of BVHs", index-augmented keys for duplicates
//   refit       : bottom-up AABB propagation with one atomic counter per internal node
//   emit        : 64 B two-child nodes (scene_dev.cuh) + TriAccel records in Morton order
#include "bvh_build.h"
#include "dev_pool.h"
#include <cub/cub.cuh>
#include <cfloat>
#include <cstring>

namespace yg {

__device__ __forceinline__ int floatFlip(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__device__ __forceinline__ float floatUnflip(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

__global__ void k_tri_setup(int n, const float *__restrict__ points, const int *__restrict__ tri_vidx, const int *__restrict__ tri_mat,
                            const unsigned *__restrict__ mat_triflags, float4 *__restrict__ tmp_v0e, float4 *__restrict__ tmp_e1i,
                            float4 *__restrict__ tmp_e2f, float4 *__restrict__ bmin, float4 *__restrict__ bmax, float *__restrict__ tri_ng,
                            int *__restrict__ scene_bound /*6 flipped ints*/)
{
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
	if(i < n)
	{
		const int ia = tri_vidx[3 * i], ib = tri_vidx[3 * i + 1], ic = tri_vidx[3 * i + 2];
		V3 a = mk(points[3 * ia], points[3 * ia + 1], points[3 * ia + 2]);
		V3 b = mk(points[3 * ib], points[3 * ib + 1], points[3 * ib + 2]);
		V3 c = mk(points[3 * ic], points[3 * ic + 1], points[3 * ic + 2]);
		V3 e1 = b - a, e2 = c - a;
		// intersection_bias_factor_ = 0.1f * MIN_RAYDIST * max(|e1|,|e2|): 0.1f * 0.00005 is a float*double product,
		// times float -> double, stored to float (triangle.h:206)
		float eps = (float)(((double)0.1f * 0.00005) * (double)fmaxf(length(e1), length(e2)));
		V3 ng = normalize(cross(e1, e2));
		tri_ng[3 * i] = ng.x; tri_ng[3 * i + 1] = ng.y; tri_ng[3 * i + 2] = ng.z;
		tmp_v0e[i] = make_float4(a.x, a.y, a.z, eps);
		tmp_e1i[i] = make_float4(e1.x, e1.y, e1.z, __int_as_float(i));
		tmp_e2f[i] = make_float4(e2.x, e2.y, e2.z, __int_as_float((int)mat_triflags[tri_mat[i]]));
		lo[0] = fminf(a.x, fminf(b.x, c.x)); hi[0] = fmaxf(a.x, fmaxf(b.x, c.x));
		lo[1] = fminf(a.y, fminf(b.y, c.y)); hi[1] = fmaxf(a.y, fmaxf(b.y, c.y));
		lo[2] = fminf(a.z, fminf(b.z, c.z)); hi[2] = fmaxf(a.z, fmaxf(b.z, c.z));
		bmin[i] = make_float4(lo[0], lo[1], lo[2], 0.f);
		bmax[i] = make_float4(hi[0], hi[1], hi[2], 0.f);
	}
	// warp reduce then one atomic per warp
	for(int k = 0; k < 3; ++k)
	{
		float l = lo[k], h = hi[k];
		for(int o = 16; o > 0; o >>= 1) { l = fminf(l, __shfl_xor_sync(0xffffffffu, l, o)); h = fmaxf(h, __shfl_xor_sync(0xffffffffu, h, o)); }
		if((threadIdx.x & 31) == 0 && l <= h) { atomicMin(&scene_bound[k], floatFlip(l)); atomicMax(&scene_bound[3 + k], floatFlip(h)); }
	}
}

__device__ __forceinline__ unsigned long long expandBits21(unsigned long long v)
{
	v &= 0x1fffffull;
	v = (v | v << 32) & 0x1f00000000ffffull;
	v = (v | v << 16) & 0x1f0000ff0000ffull;
	v = (v | v << 8) & 0x100f00f00f00f00full;
	v = (v | v << 4) & 0x10c30c30c30c30c3ull;
	v = (v | v << 2) & 0x1249249249249249ull;
	return v;
}

__global__ void k_morton(int n, const float4 *__restrict__ bmin, const float4 *__restrict__ bmax, const int *__restrict__ scene_bound,
                         unsigned long long *__restrict__ keys, int *__restrict__ vals)
{
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= n) return;
	float sl[3], sh[3];
	for(int k = 0; k < 3; ++k) { sl[k] = floatUnflip(scene_bound[k]); sh[k] = floatUnflip(scene_bound[3 + k]); }
	float4 lo = bmin[i], hi = bmax[i];
	float c[3] = {0.5f * (lo.x + hi.x), 0.5f * (lo.y + hi.y), 0.5f * (lo.z + hi.z)};
	unsigned long long q[3];
	for(int k = 0; k < 3; ++k)
	{
		float ext = sh[k] - sl[k];
		float t = ext > 0.f ? (c[k] - sl[k]) / ext : 0.f;
		t = fminf(fmaxf(t, 0.f), 1.f);
		q[k] = (unsigned long long)fminf(t * 2097152.f, 2097151.f);
	}
	keys[i] = (expandBits21(q[0]) << 2) | (expandBits21(q[1]) << 1) | expandBits21(q[2]);
	vals[i] = i;
}

// delta(i,j): common prefix length of keys i and j, augmented with the index on ties (Karras 2012 §4)
__device__ __forceinline__ int deltaFn(const unsigned long long *__restrict__ keys, int n, int i, int j)
{
	if(j < 0 || j >= n) return -1;
	unsigned long long a = keys[i], b = keys[j];
	if(a == b) return 64 + __clz(i ^ j);
	return __clzll(a ^ b);
}

// one thread per internal node i in [0, n-2]; children: internal k -> k, leaf k -> (n-1)+k in the parent array
__global__ void k_hierarchy(int n, const unsigned long long *__restrict__ keys, int *__restrict__ left, int *__restrict__ right, int *__restrict__ parent,
                            int2 *__restrict__ range)
{
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= n - 1) return;
	int d = (deltaFn(keys, n, i, i + 1) - deltaFn(keys, n, i, i - 1)) >= 0 ? 1 : -1;
	int dmin = deltaFn(keys, n, i, i - d);
	int lmax = 2;
	while(deltaFn(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
	int l = 0;
	for(int t = lmax >> 1; t >= 1; t >>= 1)
		if(deltaFn(keys, n, i, i + (l + t) * d) > dmin) l += t;
	int j = i + l * d;
	int dnode = deltaFn(keys, n, i, j);
	int s = 0;
	int t = l;
	do
	{
		t = (t + 1) >> 1;
		if(deltaFn(keys, n, i, i + (s + t) * d) > dnode) s += t;
	} while(t > 1);
	int gamma = i + s * d + min(d, 0);
	int lo = min(i, j), hi = max(i, j);
	int lc = (lo == gamma) ? (n - 1) + gamma : gamma;
	int rc = (hi == gamma + 1) ? (n - 1) + gamma + 1 : gamma + 1;
	left[i] = lc; right[i] = rc;
	range[i] = make_int2(lo, hi - lo + 1); // the Morton-ordered triangles this node covers
	parent[lc] = i; parent[rc] = i;
	if(i == 0) parent[0] = -1;
}

// bottom-up: each leaf walks to the root; the second thread to arrive at a node computes its box
__global__ void k_refit(int n, const int *__restrict__ vals, const float4 *__restrict__ bmin, const float4 *__restrict__ bmax,
                        const int *__restrict__ left, const int *__restrict__ right, const int *__restrict__ parent,
                        float4 *__restrict__ nmin, float4 *__restrict__ nmax, int *__restrict__ counters)
{
	int k = blockIdx.x * blockDim.x + threadIdx.x;
	if(k >= n) return;
	const int leaf = (n - 1) + k;
	const int t = vals[k];
	// pad each primitive box by a relative 2^-20 of its largest coordinate so that slab culling stays conservative
	// w.r.t. the (non-watertight) Möller–Trumbore arithmetic the hit test must reproduce
	float4 lo = bmin[t], hi = bmax[t];
	float m = fmaxf(fmaxf(fmaxf(fabsf(lo.x), fabsf(hi.x)), fmaxf(fabsf(lo.y), fabsf(hi.y))), fmaxf(fabsf(lo.z), fabsf(hi.z)));
	float pad = m * 9.5367431640625e-07f + 1e-30f;
	lo.x -= pad; lo.y -= pad; lo.z -= pad; hi.x += pad; hi.y += pad; hi.z += pad;
	nmin[leaf] = lo; nmax[leaf] = hi;
	__threadfence();
	int cur = parent[leaf];
	while(cur >= 0)
	{
		if(atomicAdd(&counters[cur], 1) == 0) return; // first arrival: the sibling subtree is not finished yet
		__threadfence();
		const int lc = left[cur], rc = right[cur];
		// L2 loads (bypass L1): the child boxes were written by other threads
		float4 a = __ldcg(&nmin[lc]), b = __ldcg(&nmin[rc]);
		float4 c = __ldcg(&nmax[lc]), d = __ldcg(&nmax[rc]);
		nmin[cur] = make_float4(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z), 0.f);
		nmax[cur] = make_float4(fmaxf(c.x, d.x), fmaxf(c.y, d.y), fmaxf(c.z, d.z), 0.f);
		__threadfence();
		cur = parent[cur];
	}
}

// child link: inner node index, or ~((first << 3) | (count - 1)) for a leaf of `count` consecutive TriAccel records.  A subtree
// that covers at most `leaf_max` triangles is collapsed into one leaf (its box is the subtree's box; the inner nodes below it stay
// in the array, unreferenced): fewer 64-byte node visits for a few more 48-byte triangle tests.
__device__ __forceinline__ int encodeChild(int c, int n, const int2 *__restrict__ range, int leaf_max)
{
	if(c >= n - 1) return ~(((c - (n - 1)) << 3) | 0);
	const int2 r = range[c];
	return r.y <= leaf_max ? ~((r.x << 3) | (r.y - 1)) : c;
}

__global__ void k_emit_nodes(int n, const int *__restrict__ left, const int *__restrict__ right, const float4 *__restrict__ nmin,
                             const float4 *__restrict__ nmax, const int2 *__restrict__ range, int leaf_max, BvhNode *__restrict__ nodes)
{
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= n - 1) return;
	const int lc = left[i], rc = right[i];
	float4 a = nmin[lc], b = nmax[lc], c = nmin[rc], d = nmax[rc];
	BvhNode nd;
	nd.n0 = make_float4(a.x, a.y, a.z, b.x);
	nd.n1 = make_float4(b.y, b.z, c.x, c.y);
	nd.n2 = make_float4(c.z, d.x, d.y, d.z);
	nd.n3 = make_float4(__int_as_float(encodeChild(lc, n, range, leaf_max)), __int_as_float(encodeChild(rc, n, range, leaf_max)), 0.f, 0.f);
	nodes[i] = nd;
}

__global__ void k_emit_tris(int n, const int *__restrict__ vals, const float4 *__restrict__ v0e, const float4 *__restrict__ e1i,
                            const float4 *__restrict__ e2f, TriAccel *__restrict__ out)
{
	int k = blockIdx.x * blockDim.x + threadIdx.x;
	if(k >= n) return;
	const int t = vals[k];
	TriAccel ta; ta.v0e = v0e[t]; ta.e1i = e1i[t]; ta.e2f = e2f[t];
	out[k] = ta;
}

#define CK(x) do { cudaError_t e_ = (x); if(e_ != cudaSuccess) { err = e_; goto fail; } } while(0)

cudaError_t buildBvh(const BvhBuildInput &in, BvhBuildOutput &out, cudaStream_t stream)
{
	cudaError_t err = cudaSuccess;
	const int n = in.n_tris;
	const int B = 256, G = (n + B - 1) / B;
	float4 *v0e = nullptr, *e1i = nullptr, *e2f = nullptr, *bmin = nullptr, *bmax = nullptr, *nmin = nullptr, *nmax = nullptr;
	int *scene_bound = nullptr, *vals = nullptr, *vals_s = nullptr, *left = nullptr, *right = nullptr, *parent = nullptr, *counters = nullptr;
	unsigned long long *keys = nullptr, *keys_s = nullptr;
	int2 *range = nullptr;
	const int leaf_max = in.leaf_max < 1 ? 1 : (in.leaf_max > 8 ? 8 : in.leaf_max);
	void *cub_tmp = nullptr; size_t cub_bytes = 0;
	out.nodes = nullptr; out.tris = nullptr; out.n_nodes = 0;

	CK(poolAlloc(&v0e, sizeof(float4) * n, stream)); CK(poolAlloc(&e1i, sizeof(float4) * n, stream)); CK(poolAlloc(&e2f, sizeof(float4) * n, stream));
	CK(poolAlloc(&bmin, sizeof(float4) * n, stream)); CK(poolAlloc(&bmax, sizeof(float4) * n, stream));
	CK(poolAlloc(&scene_bound, sizeof(int) * 6, stream));
	{
		int flipped[6];
		for(int k = 0; k < 6; ++k)
		{
			float f = k < 3 ? FLT_MAX : -FLT_MAX; int i; memcpy(&i, &f, 4);
			flipped[k] = i >= 0 ? i : i ^ 0x7fffffff;
		}
		CK(cudaMemcpyAsync(scene_bound, flipped, sizeof(flipped), cudaMemcpyHostToDevice, stream));
	}
	k_tri_setup<<<G, B, 0, stream>>>(n, in.points, in.tri_vidx, in.tri_mat, in.mat_triflags, v0e, e1i, e2f, bmin, bmax, in.tri_ng_out, scene_bound);
	CK(cudaGetLastError());
	CK(poolAlloc(&keys, sizeof(unsigned long long) * n, stream)); CK(poolAlloc(&keys_s, sizeof(unsigned long long) * n, stream));
	CK(poolAlloc(&vals, sizeof(int) * n, stream)); CK(poolAlloc(&vals_s, sizeof(int) * n, stream));
	k_morton<<<G, B, 0, stream>>>(n, bmin, bmax, scene_bound, keys, vals);
	CK(cudaGetLastError());
	CK(cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, keys, keys_s, vals, vals_s, n, 0, 63, stream));
	CK(poolAlloc(&cub_tmp, cub_bytes ? cub_bytes : 16, stream));
	CK(cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, keys, keys_s, vals, vals_s, n, 0, 63, stream));

	CK(poolAlloc(&out.tris, sizeof(TriAccel) * n, stream));
	k_emit_tris<<<G, B, 0, stream>>>(n, vals_s, v0e, e1i, e2f, out.tris);
	CK(cudaGetLastError());
	if(n >= 2)
	{
		CK(poolAlloc(&left, sizeof(int) * (n - 1), stream)); CK(poolAlloc(&right, sizeof(int) * (n - 1), stream));
		CK(poolAlloc(&parent, sizeof(int) * (2 * n - 1), stream)); CK(poolAlloc(&counters, sizeof(int) * (n - 1), stream));
		CK(poolAlloc(&nmin, sizeof(float4) * (2 * n - 1), stream)); CK(poolAlloc(&nmax, sizeof(float4) * (2 * n - 1), stream));
		CK(poolAlloc(&range, sizeof(int2) * (n - 1), stream));
		CK(cudaMemsetAsync(counters, 0, sizeof(int) * (n - 1), stream));
		k_hierarchy<<<(n - 1 + B - 1) / B, B, 0, stream>>>(n, keys_s, left, right, parent, range);
		CK(cudaGetLastError());
		k_refit<<<G, B, 0, stream>>>(n, vals_s, bmin, bmax, left, right, parent, nmin, nmax, counters);
		CK(cudaGetLastError());
		CK(poolAlloc(&out.nodes, sizeof(BvhNode) * (n - 1), stream));
		k_emit_nodes<<<(n - 1 + B - 1) / B, B, 0, stream>>>(n, left, right, nmin, nmax, range, leaf_max, out.nodes);
		CK(cudaGetLastError());
		out.n_nodes = n - 1;
		out.root = n <= leaf_max ? ~((0 << 3) | (n - 1)) : 0;
	}
	else
	{
		out.n_nodes = 0;
		out.root = ~0; // single leaf: first = 0, count = 1
	}
	{
		int sb[6];
		CK(cudaMemcpyAsync(sb, scene_bound, sizeof(sb), cudaMemcpyDeviceToHost, stream));
		CK(cudaStreamSynchronize(stream));
		for(int k = 0; k < 6; ++k) { int i = sb[k]; i = i >= 0 ? i : i ^ 0x7fffffff; memcpy(&out.bound[k], &i, 4); }
	}
fail:
	poolFree(v0e, stream); poolFree(e1i, stream); poolFree(e2f, stream); poolFree(bmin, stream); poolFree(bmax, stream); poolFree(nmin, stream); poolFree(nmax, stream);
	poolFree(scene_bound, stream); poolFree(vals, stream); poolFree(vals_s, stream); poolFree(left, stream); poolFree(right, stream); poolFree(parent, stream); poolFree(counters, stream);
	poolFree(keys, stream); poolFree(keys_s, stream); poolFree(cub_tmp, stream); poolFree(range, stream);
	if(err != cudaSuccess) { poolFree(out.nodes, stream); poolFree(out.tris, stream); out.nodes = nullptr; out.tris = nullptr; }
	return err;
}

} // namespace yg
