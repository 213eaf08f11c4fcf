// Wavefront state of the integrator: SoA task queues in HBM.
//
// A "task" is one ray that must be traced to its closest hit and then shaded.  Two kinds exist, mirroring the two places
// where the reference calls Scene::intersect on the hot path:
//   TASK_INTEGRATE : one call of PathIntegrator/DirectLightIntegrator::integrate() (integrator_path_tracer.cc:112-347):
//                    camera rays and the rays spawned by recursiveRaytrace (integrator_montecarlo.cc:782-1028)
//   TASK_PATH      : one vertex of the path loop of PathIntegrator::integrate (integrator_path_tracer.cc:185-295)
// Every task carries the RGB factor its result is multiplied with before it is added to the camera sample it belongs to
// (`slot`), which turns the reference's recursion tree into a flat work list.  Shadow rays are a third, lighter record:
// the contribution is computed first and added to the sample only if the any-hit trace finds the segment unoccluded.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <vector>

namespace yg {

enum { TASK_INTEGRATE = 0, TASK_PATH = 1 };

// packed task state (uint32), see packState()/fields below
//  bit 0      kind
//  bits 1-5   raylevel (0..31)
//  bits 6-11  depth of the path vertex (0..63)
//  bits 12-19 ray_division (1..255)
//  bit 22     caustic
//  bits 23-27 additional_depth (0..31)
struct TaskQueue
{
	float4 *ray_o;   // from.xyz, tmin
	float4 *ray_d;   // dir.xyz, tmax (<0 = infinite)
	float4 *hit;     // t, u, v, bits(tri)     written by k_trace
	float4 *wgt;     // rgb factor from this task's result up to the camera sample
	float4 *thr;     // TASK_PATH: throughput of the path alone (Russian roulette uses it, integrator_path_tracer.cc:285)
	uint4 *st;       // slot, offs (path) , packed, ray_offset
	float2 *dc;      // dc_1, dc_2
};

struct ShadowQueue
{
	float4 *ray_o;   // from.xyz (NOT yet shifted), tmin
	float4 *ray_d;   // dir.xyz, tmax
	float4 *contrib; // rgb contribution if unoccluded, bits(slot)
};

enum { CNT_RAYS_CAMERA = 0, CNT_RAYS_EXTENSION, CNT_RAYS_SHADOW, CNT_RAYS_SHADOW_TRANSP, CNT_RAYS_LIGHT_PROBE,
       CNT_CAMERA_SAMPLES, CNT_NODES, CNT_TRIS, CNT_NODES_SHADOW, CNT_TRIS_SHADOW, CNT_NONFINITE, CNT_N };

#define YG_MAX_ITERS 96
#define YG_N_MAT_TYPES 11  /* yafgpu_material_type values 0..8 + YG_MAT_SD_LAMBERT + YG_MAT_NODES */
#define YG_MAX_BUCKETS 8200 /* 2 task kinds x up to 4096 material buckets + miss */
#define YG_SHARED_BUCKETS 2048 /* up to this many buckets the histogram and the scatter ranks are staged per block in shared memory */

struct WaveCtl // lives in device memory, zeroed at wave start
{
	int n_tasks[YG_MAX_ITERS + 2];   // number of tasks queued for iteration i
	int n_shadow[YG_MAX_ITERS + 2];  // number of shadow rays queued by iteration i
	int fetch_trace[YG_MAX_ITERS + 2];  // next unclaimed task of iteration i (ray regeneration in k_trace)
	int fetch_shadow[YG_MAX_ITERS + 2]; // next unclaimed shadow ray of iteration i (k_shadow)
	int overflow;                     // a queue ran out of capacity: the wave is discarded and retried smaller
	int pad;
};

struct Wavefront
{
	TaskQueue q[2];
	ShadowQueue sq;
	float4 *acc = nullptr;        // per camera sample: rgb sum, alpha
	uint32_t *pixels = nullptr;   // pass pixel list: x | y << 16
	size_t pixels_cap = 0, pixels_n = 0;
	std::vector<int> pixels_key; // parameters the cached pixel list was built for
	WaveCtl *ctl = nullptr;
	int *key = nullptr, *perm = nullptr; // per task: sort bucket, and the bucket-ordered permutation read by the shading kernel
	int *hist = nullptr;                 // [2][YG_MAX_BUCKETS]: bucket counts, bucket offsets
	WaveCtl *h_ctl = nullptr;     // pinned mirror
	int wave_cap = 0;             // camera samples per wave
	int task_cap = 0;             // tasks per queue
	int shadow_cap = 0;           // shadow rays per iteration
	bool allocated = false;
};

} // namespace yg
