// yafgpu context: owns all device memory of one GPU (internal header, not part of the C ABI).
#pragma once
#include "../../include/yafgpu.h"
#include "scene_dev.cuh"
#include "wavefront.h"
#include <string>
#include <vector>

#ifndef YG_DEFAULT_REGEN
#define YG_DEFAULT_REGEN 1
#endif
#ifndef YG_DEFAULT_LEAF
#define YG_DEFAULT_LEAF 2 /* triangles per leaf link of the wide BVH unless yafgpu_set_option("leaf_size") says otherwise */
#endif

struct yafgpu_ctx
{
	int device = 0;
	unsigned flags = 0;
	cudaStream_t stream = nullptr;
	std::string err;
	bool uploaded = false;

	// scene arrays (device)
	float *d_points = nullptr, *d_normals = nullptr, *d_uv = nullptr, *d_tri_ng = nullptr;
	int *d_tri_vidx = nullptr, *d_tri_mat = nullptr;
	unsigned *d_mat_triflags = nullptr;
	int type_b0[16] = {0}, type_b1[16] = {-1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1}; // per material type: bucket range, b1 < b0 = type absent
	int *d_mat_bucket = nullptr; int n_mat_buckets = 1; // material -> sort bucket (rank in (type, id) order; one bucket per type above 4096 materials)
	yg::TriShade *d_tris = nullptr;
	yg::ObjDev *d_objects = nullptr;
	yg::MatDev *d_mats = nullptr;
	yg::LightDev *d_lights = nullptr;
	float *d_light_tables = nullptr;
	yg::Accel *d_light_accels = nullptr;
	std::vector<void *> light_bvh_ptrs; // node/triangle arrays of the mesh-light accelerators
	float *d_bg_tex = nullptr, *d_cie = nullptr;
	yg::NoiseTables *d_noise = nullptr;          // lattice tables of the noise bases (only with clouds / marble / wood textures)
	yg::TexDev *d_textures = nullptr; yg::NodeDev *d_shader_nodes = nullptr; // image textures and shader nodes (nodes.cuh)
	std::vector<void *> tex_ptrs;                                            // texel arrays of d_textures
	yg::FaureTables *d_faure = nullptr;
	void *d_nodes = nullptr;
	yg::TriAccel *d_acc_tris = nullptr;
	size_t n_tris = 0, n_points = 0, n_normals = 0, n_uv = 0, n_objects = 0, n_mats = 0, n_lights = 0;

	yg::SceneDev scene;   // by-value kernel argument
	yg::FilmDev film;     // by-value kernel argument
	float scene_bound[6] = {0, 0, 0, 0, 0, 0};
	std::vector<yg::MatDev> h_mats;
	std::vector<yg::LightDev> h_lights;
	int tile_size = 32;
	int wave_cur_cap = 0;                             // wave size after queue overflows on the current scene (0 = none yet); reset by every upload
	int wave_cap_override = 0, task_fan_override = 0; // yafgpu_set_option

	// wavefront state
	yg::Wavefront wf;

	// stats
	yafgpu_stats stats;
	int leaf_max = YG_DEFAULT_LEAF;                         // yafgpu_set_option("leaf_size"): BVH leaf collapse threshold used by the NEXT upload
	int ray_regeneration = -1;                // yafgpu_set_option("ray_regeneration"): -1 = library default (YG_DEFAULT_REGEN), 0 none, 1 both kernels, 2 closest-hit only, 3 shadow only
	int regen_threshold = 20;                 // yafgpu_set_option("regen_threshold"): refill a warp as soon as fewer lanes than this are busy
	int *d_fetch = nullptr;                   // ray-regeneration counter of the batch trace kernel
	int *d_fault = nullptr, *h_fault = nullptr; // traversal-stack fault word (Accel::fault) and its pinned mirror
	std::vector<uint8_t> aa_mask; bool aa_mask_valid = false; // resampling mask of the last yafgpu_next_pass (empty = every pixel), used by passes with adaptive = 2
	int adaptive_iters = 1;                   // yafgpu_set_option("adaptive_iterations"): 0 = always launch the schedule's maximum number of iterations (A/B)
	int iters_hint = 0;                       // wavefront iterations the previous wave of this scene needed + 1 (0 = unknown: launch the schedule's maximum)
	int keep_nonfinite = 0;                   // yafgpu_set_option("keep_nonfinite_samples"): splat NaN/Inf camera samples as the reference does instead of dropping them
	int stack_limit = 0;                      // yafgpu_set_option("debug_stack_limit"): 0 = YG_STACK
	unsigned long long *d_counters = nullptr; // device-side counters (see wavefront.h CNT_*)
	cudaEvent_t ev0 = nullptr, ev1 = nullptr;
	cudaStream_t side[2] = {nullptr, nullptr};   // the per-type shading kernels of one iteration run side by side (integrator.cu)
	cudaEvent_t ev_fork = nullptr, ev_join[2] = {nullptr, nullptr};
	int shade_streams = 3;                       // yafgpu_set_option("shade_streams", 1..3)
	std::vector<cudaEvent_t> stage_events; // YAFGPU_FLAG_TIME_STAGES
};

namespace yg {
int setErr(yafgpu_ctx *ctx, int code, const std::string &msg);
int cudaErr(yafgpu_ctx *ctx, cudaError_t e, const char *what);
int renderPass(yafgpu_ctx *ctx, const yafgpu_pass_desc *pass); // integrator.cu
int nextPassMask(yafgpu_ctx *ctx, const float *film, const yafgpu_aa_noise_params *prm, std::vector<uint8_t> &mask, int64_t *n_resample); // api.cu
void freeWavefront(yafgpu_ctx *ctx);
int faultCheck(yafgpu_ctx *ctx, bool fetch); // api.cu: YAFGPU_ERR_CUDA when a traversal stack overflowed since the last check (fetch: copy the word first)
}

#define YG_CUDA(ctx, call) do { cudaError_t e__ = (call); if(e__ != cudaSuccess) return yg::cudaErr(ctx, e__, #call); } while(0)
