// HBM layout of the flattened scene.  Plain-old-data views handed to every kernel by value.
#pragma once
#include "../../include/yafgpu.h"
#include "dev_math.cuh"
#include "dev_qmc.cuh"

namespace yg {

// ---- accelerator ---------------------------------------------------------------------------------------------------
// Triangle record used by traversal: 48 B, three float4, stored in BVH (Morton) order.
//   v0e  = (a.x, a.y, a.z, eps)           a = vertex pa_, eps = intersection_bias_factor_ (triangle.h:197-207)
//   e1i  = (edge_1.xyz, bits of global triangle id)
//   e2f  = (edge_2.xyz, bits of flags)    flags: TRI_VIS_CAMERA | TRI_VIS_SHADOW | TRI_TRANSPARENT
// Full float32 precision is mandatory: Möller–Trumbore must reproduce Triangle::intersect (triangle.h:223-259) exactly.
struct TriAccel { float4 v0e, e1i, e2f; };
enum { TRI_VIS_CAMERA = 1,  // visibility in {Normal, VisibleNoShadows}      kdtree_triangle.cc:786
       TRI_VIS_SHADOW = 2,  // visibility in {Normal, InvisibleShadowsOnly}  kdtree_triangle.cc:938
       TRI_TRANSPARENT = 4  // Material::isTransparent()                     kdtree_triangle.cc:1106
};

// W-wide BVH node, W = 4 (128 B = one cache line) or W = 8 (256 B), float32 child planes, laid out in 32-byte chunks so that the
// traversal fetches a node with 256-bit loads (LDG.E.256: one L1 wavefront per chunk and lane instead of two 128-bit ones — the
// traversal kernels sit at 80-93 % of the L1 pipe, profiles/r2_ncu_summary.txt).  With V = W / 4:
//   chunk a * V + v (a = 0, 1, 2 = x, y, z): [ lo.a of children 4v..4v+3 | hi.a of children 4v..4v+3 ]
//   chunk 3 * V + v                        : [ links of children 4v..4v+3 | child count, padding ]
//   link >= 0 inner node index; < 0 leaf, ~link = (first << 3) | (count - 1) into the TriAccel array; 0x80000000 = no child (its box
//   is inverted, lo = +FLT_MAX > hi, so the box test alone rejects it)
// The TriAccel records stay full float32: the boxes only cull, Triangle::intersect decides.
#define YG_NO_CHILD ((int)0x80000000)

struct Accel
{
	const uint4 *nodes;
	const TriAccel *tris;
	int n_nodes, n_tris;
	int root;          // encoded like a link (a scene of up to leaf_max triangles has a leaf root)
	int width;         // 4 or 8
	float bmin[3], bmax[3]; // tree bound grown by 0.1 % as TriKdTree does (kdtree_triangle.cc:112-116)
	int stack_limit;   // traversal stack entries in use (YG_STACK; lower only through the debug option that tests the fault path)
	int *fault;        // device word raised by a traversal whose stack overflowed (traverse.cuh stackFault); owned by the context
};

// ---- shading data (global triangle order) ---------------------------------------------------------------------------
struct __align__(16) TriShade // 64 B = four 16 B loads
{
	int v[3];     // indices into points
	int n[3];     // indices into normals or -1
	int uv[3];    // indices into uv or -1
	int mat;      // material index
	int obj;      // object index
	float ng[3];  // geometric normal, Triangle::recNormal (triangle.h:274-281)
	int pad[2];
};

struct ObjDev { int has_orco, has_uv, smooth, light; };

// internal material type (never crosses the C ABI): shinydiffuse with the diffuse-reflection component only, see materials.cuh
#define YG_MAT_SD_LAMBERT 9
// kernel class (never a MatDev::type) of every material that carries shader nodes: generic type dispatch + node evaluation, nodes.cuh
#define YG_MAT_NODES 10

// Material record after the reference constructors' preprocessing (one per yafgpu_material).
struct MatDev
{
	int type, visibility, receive_shadows, flat;
	int additional_depth, transp_bias_mult;
	float transp_bias_factor, sampling_factor;
	unsigned bsdf_flags;   // Material::bsdf_flags_
	// shinydiffuse (material_shiny_diffuse.cc:25-92,599-688)
	float diffuse_col[3], mirror_col[3], emit_col[3];
	float diffuse_strength, mirror_strength, transparency_strength, translucency_strength, emit_strength;
	float transmit_filter_strength;
	float ior, ior_squared;
	int has_fresnel, is_mirror, is_transparent, is_translucent, is_diffuse;
	int n_bsdf; unsigned c_flags[4]; int c_index[4];
	int use_oren_nayar; float oren_a, oren_b;
	// glass / rough glass (material_glass.cc:32-49, material_rough_glass.cc)
	float filter_col[3], spec_refl_col[3];
	unsigned tm_flags; int fake_shadow, disperse; float cauchy_a, cauchy_b; float alpha_rough;
	int absorb; float beer_sigma_a[3];
	// glossy / coated glossy
	float gloss_col[3], diff_col[3], gl_mirror_col[3];
	float gl_diffuse, gl_glossy, gl_exponent, exp_u, exp_v, gl_mirror_strength, p_diffuse;
	int as_diffuse, anisotropic, with_diffuse, gl_oren_nayar; float gl_oren_a, gl_oren_b;
	// blend
	int mat1, mat2; float blend_val;
	// simple (light_mat / mirror)
	float simple_col[3]; int double_sided;
	// shader nodes (nodes.cuh): this material's nodes are SceneDev::nodes[node_first .. node_first + node_count) in evaluation order
	// (stack index = position); bit i of color_mask / bump_mask = node i is in the colour / bump node list; slot[] = stack index of the
	// node bound to each "<slot>_shader" parameter, -1 = none.  Filled per hit from the node results (evalMaterialNodes):
	int node_first, node_count; unsigned color_mask, bump_mask; int slot[YAFGPU_N_SLOTS];
	unsigned obj_flags; int is_mask;     // Material::getFlags() of the object the integrator holds: equal to bsdf_flags except under a mask_mat, where
	                                     // bsdf_flags is the chosen child's (what initBsdf returns) and obj_flags the union of both children
	float diffuse_refl_mul;              // shinydiffuse "diffuse_refl_shader" (1 without it: the reference then does not multiply)
	int oren_tex; float oren_sigma;      // "sigma_oren_shader": Oren-Nayar coefficients from the textured sigma, in double (material_shiny_diffuse.cc:231-236)
};

#define YG_MAX_NODES 16 /* nodes per material (per-thread node stack: 20 B each) */

// Lattice tables of the reference's noise bases in the byte layout of yafgpu_scene_desc.noise_tables (noise_generator.cc:25-320)
struct NoiseTables
{
	float hashpntf[768];      // Voronoi feature points / cell colours
	float hashvectf[768];     // Blender noise gradients
	float stdp_g[514 * 3];    // standard Perlin gradients
	unsigned char hash[512];  // the permutation shared by new Perlin, Blender noise and Voronoi
	unsigned char stdp_p[516];// standard Perlin permutation (514 used)
};
static_assert(sizeof(NoiseTables) == YAFGPU_NOISE_TABLES_BYTES, "noise table blob layout");

// ImageTexture after its factory (texture_image.cc:560-720); texels as ImageHandler::getPixel returns them, RGBA float
struct TexDev
{
	const float4 *tex; int w, h;
	int interp, clip, xrepeat, yrepeat, mirror_x, mirror_y, rot90, cropx, cropy;
	float cropminx, cropminy, cropmaxx, cropmaxy;
	int checker_even, checker_odd; float checker_dist;
	int normalmap, color_space; float gamma;
	int type, blend_type, flip;  // yafgpu_texture_type; blend: progression type, use_flip_axis
	int adj_set, adj_clamp; float adj_intensity, adj_contrast, adj_saturation, adj_hue, adj_mult[3]; // adj_hue already / 60 (texture.h:154)
	// clouds / marble / wood (texture_basic.cc:96-343)
	const NoiseTables *noise; int noise_type, depth, hard, bias, shape, rings; float size, turb, sharpness, color1[3], color2[3];
	// voronoi (texture_basic.cc:391-453): weights, their magnitudes, isc / sum of magnitudes; musgrave (:509-551); distorted noise (:602-633)
	float vw[4], vaw[4], iscale, mk_exp; int dist_metric, color_type;
	int mus_type; float mus_h, mus_lac, mus_oct, mus_offset, mus_gain;
	float distort; int noise_type2;
	// colour ramp (color_ramp.cc): n_ramp items of r, g, b, a, position
	const float *ramp; int n_ramp, ramp_mode, ramp_interp, ramp_hue;
};

// One shader node after its factory + configInputs; in1/in2/fac are stack indices (-1 = constant input)
struct NodeDev
{
	int type, in1, in2, fac, mode;
	// texture mapper after setup() (shader_node_basic.cc:33-55): offset doubled, bump_str scaled, d_u/d_v = 1 / resolution
	int tex, coords, proj, map_x, map_y, map_z, do_scalar;
	float scale[3], offset[3], bump_str, d_u, d_v;
	int discrete;                    // Texture::discrete(): image textures; procedural ones take the 3-D finite-difference bump (shader_node_basic.cc:240,320-338)
	float mtx[12];
	float col[4], col2[4], f1, f2;   // value: colour, scalar; mix: color1, color2, cfactor; layer: def_col, upper_col, def_val, upper_val
	float colfac, valfac; unsigned texflag; int do_color, do_scalar_l, color_input, use_alpha;
};

// Light record after constructor preprocessing.
struct LightDev
{
	int type, enabled, cast_shadows, photon_only, samples, object, double_sided, abs_intersect;
	float color[3];     // post-constructor colour (e.g. area: col*power*pi)
	float pos[3];       // point position / sun direction / area corner
	float to_x[3], to_y[3], fnormal[3], normal[3]; // area
	float c2[3], c3[3], c4[3];                     // area corners
	float area, inv_area;
	// sun (light_sun.cc:29-55)
	float du[3], dv[3]; float cos_angle, pdf, invpdf; float col_pdf[3];
	// mesh / background: offsets into LightTables
	int cdf_offset, cdf_count; float cdf_integral;
	int tri_first, tri_count;
	int accel;          // mesh light: index into SceneDev::light_accels (its own BVH, as MeshLight builds its own kd-tree), -1 = none
	float clamp_intersect; int shoot_caustic;
	float power;        // background portal: colour = Background::eval(dir) * power_
	int dirac, can_intersect;   // Light::diracLight() / canIntersect() (the spot light's depend on "soft_shadows")
	int ies_nv, ies_nh, ies_type; float ies_max_rad; // ies: IesData (util_ies.h:33-56); tables at cdf_offset: vert angles [nv] +INF, hor angles [nh] -INF, rad_map [nh][nv]
	// directional (pos = position_, fnormal = direction_, area = radius_, double_sided = !infinite_), spot (pos, fnormal = ndir_, normal = dir_,
	// du/dv, cos_angle = cos_end_, pdf = cos_start_, invpdf = icos_diff_, area = shadow_fuzzy_), sphere (pos = center, area = radius^2,
	// inv_area = radius^2 * 1.000003815)
};

struct CameraDev
{
	float pos[3], vright[3], vup[3], vto[3], dof_rt[3], dof_up[3];
	float near_n[3], near_p[3], far_p[3];
	float aperture, dof_distance;
	int resx, resy, bokeh_type, bokeh_bias, n_ls;
	float ls[20];
	int type;                    // yafgpu_camera_type; architect and perspective differ only in the host-side axes
	float aspect_ratio;          // Camera::aspect_ratio_ = aspect * resy / resx (camera.cc:47)
	float cam_to[3];             // orthographic / angular / equirectangular: vto_ = cam_z_ (vright/vup hold their own vright_/vup_)
	float max_radius, focal_length; int circular, projection; // angular (camera_angular.cc:28-42)
	float ortho_scale;           // OrthographicCamera::scale_ (screenproject)
	float axis_x[3], axis_y[3], axis_z[3], focal;              // Camera::cam_x_/cam_y_/cam_z_ (unit axes, camera.cc:46-66) and PerspectiveCamera::focal_distance_
};

struct BackgroundDev
{
	int type; float color[3]; // constant: col*power
	int has_ibl, shoots_caustic;
	// texture: background_texture.cc:32-88.  rotation = 2 * rot / 360, sin_r/cos_r = fSin/fCos(pi * rotation); tex = RGB float texels
	// as ImageHandler::getPixel(x, y) returns them, row y at tex + 3 * y * tex_w (device pointer in kernels, host pointer while the
	// background-light tables are built)
	int mapping_angular; float rotation, sin_r, cos_r, power; const float *tex; int tex_w, tex_h, tex_interp;
	float gzenith[3], ghoriz[3], szenith[3], shoriz[3]; // gradient: colours * power (background_gradient.cc:88)
};

struct IntegratorDev
{
	int type, r_depth, tr_shad, s_depth, n_paths, max_bounces, rr_min_bounces, no_recursive, trace_caustics;
	int transp_bg, transp_refr_bg, do_ao, ao_samples; float ao_dist, ao_col[3];
	float shadow_bias; int shadow_bias_auto; float ray_min_dist; int ray_min_dist_auto;
	float aa_light_sample_multiplier, aa_indirect_sample_multiplier, aa_clamp_indirect;
};

struct FilmDev
{
	float *pix;         // w*h*5: r,g,b,a,weight
	int w, h, cx0, cy0, cx1, cy1;
	float filterw, table_scale, clamp_samples;
	int premult, aa_passes;
	int transp_bg;      // "bg_transp": a camera sample whose primary ray left the scene has alpha 0 (acc[slot].w carries the mark)
	unsigned base_sampling_offset;
	float table[256];   // 16x16 filter table, imagefilm.cc:167-174
};

struct SceneDev
{
	Accel accel;
	const float *points, *normals, *uv;
	const TriShade *tris;
	const ObjDev *objects;
	const MatDev *materials;
	const LightDev *lights; int n_lights;
	const float *light_tables;
	const Accel *light_accels;
	const FaureTables *faure;
	const TexDev *textures;
	const NodeDev *nodes;
	const float *cie_xyz; int n_cie;   // CIE 1931 colour matching functions, rows of x,y,z from 360 nm in 1 nm steps (dispersion only)
	CameraDev cam;
	BackgroundDev bg;
	IntegratorDev integ;
};

} // namespace yg
