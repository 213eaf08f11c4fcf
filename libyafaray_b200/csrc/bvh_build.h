// K1 interface: GPU LBVH build (bvh_build.cu).
#pragma once
#include "scene_dev.cuh"

namespace yg {

struct BvhBuildInput
{
	int n_tris;
	const float *points;          // device
	const int *tri_vidx;          // device, 3 per triangle
	const int *tri_mat;           // device
	const unsigned *mat_triflags; // device, per material TRI_* flags
	float *tri_ng_out;            // device, 3 per triangle: geometric normals (written)
	int leaf_max = 1;             // subtrees of at most this many triangles become one leaf (1..8)
	int width = 8;                // children per node: 4 (64 B nodes) or 8 (96 B nodes)
};

struct BvhBuildOutput
{
	void *nodes; int n_nodes;     // quantised wide nodes (scene_dev.cuh)
	int node_bytes, levels;
	TriAccel *tris;
	int root;
	float bound[6]; // exact scene bound: min xyz, max xyz
};

cudaError_t buildBvh(const BvhBuildInput &in, BvhBuildOutput &out, cudaStream_t stream);

} // namespace yg
