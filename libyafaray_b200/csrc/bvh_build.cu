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
//   collapse    : the binary hierarchy is collapsed top-down, level by level, into W-wide nodes (W = 4 or 8): a wide node starts from
//                 the two children of its binary node and repeatedly opens the child with the largest surface area until it holds W
//                 children (the SAH-greedy collapse of Wald et al. 2008 / Ylitie et al. 2017); subtrees of at most `leaf_max`
//                 triangles become leaf links
//   emit        : W-wide nodes with float32 child planes in structure-of-arrays form (scene_dev.cuh) + TriAccel records in Morton order.
//                 (An 8-bit quantised node was built and measured first: the traversal kernel is bound by instruction issue — 68 %
//                 issue-slot utilisation at 10.5 of 32 lanes, DRAM at 4 % — so the byte->float conversions cost more than the smaller
//                 nodes saved: profiles/r2b_*.)
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
	// cubic cells: every axis is scaled by the LARGEST extent, so a flat scene (a terrain) spends its code bits on the two long axes
	// instead of splitting its thin axis as often as the others
	const float ext = fmaxf(fmaxf(sh[0] - sl[0], sh[1] - sl[1]), sh[2] - sl[2]);
	for(int k = 0; k < 3; ++k)
	{
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

// ---- collapse to W-wide quantised nodes ------------------------------------------------------------------------------------------
// A binary node id is: k in [0, n-1) internal node k, (n-1)+k leaf k (triangle at Morton position k).
struct CollapseItem { int node2; int slot; };   // binary node to expand -> wide node index to write

__device__ __forceinline__ float boxArea(const float4 lo, const float4 hi)
{
	const float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
	return dx * dy + dy * dz + dz * dx;
}

// triangles under binary node c (1 for a leaf)
__device__ __forceinline__ int trisUnder(int c, int n, const int2 *__restrict__ range) { return c >= n - 1 ? 1 : range[c].y; }

template<int W>
__global__ void k_collapse(int n, int n_items, const CollapseItem *__restrict__ items, const int *__restrict__ left, const int *__restrict__ right,
                           const float4 *__restrict__ nmin, const float4 *__restrict__ nmax, const int2 *__restrict__ range, int leaf_max,
                           uint4 *__restrict__ nodes, CollapseItem *__restrict__ next_items, int *__restrict__ counters /* [0] wide nodes, [1] next queue length */)
{
	const int it = blockIdx.x * blockDim.x + threadIdx.x;
	if(it >= n_items) return;
	const CollapseItem item = items[it];
	int ch[W]; float area[W];   // area < 0: not openable (leaf link)
	int cnt = 2;
	ch[0] = left[item.node2]; ch[1] = right[item.node2];
#pragma unroll
	for(int k = 0; k < W; ++k)
	{
		if(k < 2) area[k] = trisUnder(ch[k], n, range) > leaf_max ? boxArea(nmin[ch[k]], nmax[ch[k]]) : -1.f;
		else { ch[k] = -1; area[k] = -1.f; }
	}
	while(cnt < W)
	{
		int best = -1; float ba = -1.f;
#pragma unroll
		for(int k = 0; k < W; ++k) if(k < cnt && area[k] > ba) { ba = area[k]; best = k; }   // area >= 0 only for openable children
		if(best < 0) break;
		int c = -1;
#pragma unroll
		for(int k = 0; k < W; ++k) if(k == best) c = ch[k];
		const int l = left[c], r = right[c];
		const float al = trisUnder(l, n, range) > leaf_max ? boxArea(nmin[l], nmax[l]) : -1.f;
		const float ar = trisUnder(r, n, range) > leaf_max ? boxArea(nmin[r], nmax[r]) : -1.f;
#pragma unroll
		for(int k = 0; k < W; ++k)
		{
			if(k == best) { ch[k] = l; area[k] = al; }
			if(k == cnt) { ch[k] = r; area[k] = ar; }
		}
		++cnt;
	}
	// inner children get consecutive wide-node slots
	int n_inner = 0;
#pragma unroll
	for(int k = 0; k < W; ++k) if(k < cnt && area[k] >= 0.f) ++n_inner;
	int slot0 = 0, q0 = 0;
	if(n_inner) { slot0 = atomicAdd(&counters[0], n_inner); q0 = atomicAdd(&counters[1], n_inner); }
	// node (scene_dev.cuh): per axis the lo and hi planes of W children, then W links, 128 B (W = 4) / 256 B (W = 8)
	float pl[6][W]; int link[W];
	int ii = 0;
#pragma unroll
	for(int k = 0; k < W; ++k)
	{
		if(k >= cnt)
		{	// no child: inverted box, never hit
			for(int a = 0; a < 3; ++a) { pl[a][k] = FLT_MAX; pl[3 + a][k] = -FLT_MAX; }
			link[k] = (int)0x80000000;
			continue;
		}
		const int c = ch[k];
		const float4 lo = nmin[c], hi = nmax[c];
		pl[0][k] = lo.x; pl[1][k] = lo.y; pl[2][k] = lo.z; pl[3][k] = hi.x; pl[4][k] = hi.y; pl[5][k] = hi.z;
		if(area[k] >= 0.f)
		{
			link[k] = slot0 + ii;
			next_items[q0 + ii] = CollapseItem{c, slot0 + ii};
			++ii;
		}
		else if(c >= n - 1) link[k] = ~(((c - (n - 1)) << 3) | 0);
		else { const int2 r = range[c]; link[k] = ~((r.x << 3) | (r.y - 1)); }
	}
	// 32-byte chunks (one 256-bit load each): per axis [lo of children 4v..4v+3 | hi of the same], then [links | child count, padding]
	float4 *out = (float4 *)(nodes + (size_t)item.slot * (W == 4 ? 8 : 16));
#pragma unroll
	for(int a = 0; a < 3; ++a)
#pragma unroll
		for(int v = 0; v < W / 4; ++v)
		{
			out[2 * (a * (W / 4) + v)] = make_float4(pl[a][4 * v], pl[a][4 * v + 1], pl[a][4 * v + 2], pl[a][4 * v + 3]);
			out[2 * (a * (W / 4) + v) + 1] = make_float4(pl[3 + a][4 * v], pl[3 + a][4 * v + 1], pl[3 + a][4 * v + 2], pl[3 + a][4 * v + 3]);
		}
#pragma unroll
	for(int v = 0; v < W / 4; ++v)
	{
		out[2 * (3 * (W / 4) + v)] = make_float4(__int_as_float(link[4 * v]), __int_as_float(link[4 * v + 1]), __int_as_float(link[4 * v + 2]), __int_as_float(link[4 * v + 3]));
		out[2 * (3 * (W / 4) + v) + 1] = make_float4(__int_as_float(cnt), 0.f, 0.f, 0.f);
	}
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

__global__ void k_pad_boxes(int n_boxes, float4 *__restrict__ nmin, float4 *__restrict__ nmax, float pad)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= n_boxes) return;
	float4 a = nmin[i], b = nmax[i];
	a.x -= pad; a.y -= pad; a.z -= pad; b.x += pad; b.y += pad; b.z += pad;
	nmin[i] = a; nmax[i] = b;
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
	CollapseItem *items[2] = {nullptr, nullptr}; int *ccount = nullptr;
	const int leaf_max = in.leaf_max < 1 ? 1 : (in.leaf_max > 8 ? 8 : in.leaf_max);
	void *cub_tmp = nullptr; size_t cub_bytes = 0;
	out.nodes = nullptr; out.tris = nullptr; out.n_nodes = 0; out.node_bytes = in.width == 4 ? 128 : 256; out.levels = 0;

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
		// pad every box by 2^-19 of the scene's largest coordinate: the traversal tests the planes in t-space with fused arithmetic
		// whose rounding differs from Moeller-Trumbore's by a few ulps of the coordinates involved
		{
			int sb[6];
			CK(cudaMemcpyAsync(sb, scene_bound, sizeof(sb), cudaMemcpyDeviceToHost, stream));
			CK(cudaStreamSynchronize(stream));
			float mx = 0.f;
			for(int k = 0; k < 6; ++k) { int i = sb[k]; i = i >= 0 ? i : i ^ 0x7fffffff; float f; memcpy(&f, &i, 4); mx = fmaxf(mx, fabsf(f)); }
			const float pad = mx * 1.9073486328125e-06f + 1e-30f;
			k_pad_boxes<<<(2 * n - 1 + B - 1) / B, B, 0, stream>>>(2 * n - 1, nmin, nmax, pad);
			CK(cudaGetLastError());
		}
		const int W = in.width == 4 ? 4 : 8;
		const size_t node_u4 = W == 4 ? 8 : 16;
		out.node_bytes = (int)(node_u4 * sizeof(uint4));
		if(n <= leaf_max) { out.n_nodes = 0; out.root = ~((0 << 3) | (n - 1)); }
		else
		{
			CK(poolAlloc(&out.nodes, sizeof(uint4) * node_u4 * (size_t)(n - 1), stream));
			CK(poolAlloc(&items[0], sizeof(CollapseItem) * (size_t)n, stream)); CK(poolAlloc(&items[1], sizeof(CollapseItem) * (size_t)n, stream));
			CK(poolAlloc(&ccount, sizeof(int) * 2, stream));
			const CollapseItem root_item = {0, 0};
			int hc[2] = {1, 0};
			CK(cudaMemcpyAsync(items[0], &root_item, sizeof(root_item), cudaMemcpyHostToDevice, stream));
			CK(cudaMemcpyAsync(ccount, hc, sizeof(hc), cudaMemcpyHostToDevice, stream));
			int n_items = 1, cur = 0, levels = 0;
			while(n_items > 0)
			{
				const int g = (n_items + 127) / 128;
				if(W == 4) k_collapse<4><<<g, 128, 0, stream>>>(n, n_items, items[cur], left, right, nmin, nmax, range, leaf_max, (uint4 *)out.nodes, items[cur ^ 1], ccount);
				else k_collapse<8><<<g, 128, 0, stream>>>(n, n_items, items[cur], left, right, nmin, nmax, range, leaf_max, (uint4 *)out.nodes, items[cur ^ 1], ccount);
				CK(cudaGetLastError());
				CK(cudaMemcpyAsync(hc, ccount, sizeof(hc), cudaMemcpyDeviceToHost, stream));
				CK(cudaMemsetAsync(ccount + 1, 0, sizeof(int), stream));
				CK(cudaStreamSynchronize(stream));
				n_items = hc[1]; cur ^= 1; ++levels;
			}
			out.n_nodes = hc[0]; out.root = 0; out.levels = levels;
		}
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
	poolFree(items[0], stream); poolFree(items[1], stream); poolFree(ccount, stream);
	if(err != cudaSuccess) { poolFree(out.nodes, stream); poolFree(out.tris, stream); out.nodes = nullptr; out.tris = nullptr; }
	return err;
}

} // namespace yg
