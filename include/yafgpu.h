/* yafgpu.h — C ABI of the B200-native path-tracing hot path for libYafaRay.
 *
 * This is the drop-in boundary.  Everything a
 * reference-side adaptor needs in order to replace the CPU hot path
 *
 *   Scene::update() accelerator build      /root/reference/src/common/scene.cc:788-831
 *   SurfaceIntegrator::render()            /root/reference/include/integrator/integrator.h:73-85
 *     = TiledIntegrator::render/renderPass /root/reference/src/integrator/integrator_tiled.cc:116-345
 *   ImageFilm::addSample accumulation      /root/reference/src/common/imagefilm.cc:925-1015
 *
 * is exported here as plain C: POD structs, pointers and sizes only; no C++
 * types, no exceptions, no torch types.  The caller keeps ownership of every
 * pointer it passes in; the library copies what it needs before returning.
 * Every function returns 0 on success and a non-zero yafgpu_status otherwise;
 * yafgpu_last_error() gives the message.  There is NO CPU fallback: if no
 * CUDA device is usable, yafgpu_create() fails.
 *
 * Descriptors are PARAMETER-level for cameras, lights, materials, film and
 * integrator: the fields are the values the reference's factories read from
 * their ParamMap (after defaults), cited per struct.  Derived constants (camera
 * axes, light areas, CDFs, filter tables) are computed inside the library the
 * way the reference constructors compute them.  Geometry is flattened: triangles
 * in the reference's GLOBAL TRIANGLE ORDER (scene.cc:807-817: ascending object
 * id, each mesh in addTriangle order), which is also the id space of every hit
 * record this API returns.
 */
#ifndef YAFGPU_H
#define YAFGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define YAFGPU_ABI_VERSION 5

typedef struct yafgpu_ctx yafgpu_ctx;

typedef enum yafgpu_status
{
	YAFGPU_OK = 0,
	YAFGPU_ERR_INVALID = 1,     /* bad argument / malformed descriptor */
	YAFGPU_ERR_UNSUPPORTED = 2, /* feature of the reference not implemented on the device (refused, never approximated) */
	YAFGPU_ERR_CUDA = 3,        /* CUDA runtime error */
	YAFGPU_ERR_STATE = 4,       /* call order violated (e.g. render before upload) */
	YAFGPU_ERR_NOMEM = 5
} yafgpu_status;

/* ---- geometry -------------------------------------------------------------------------- */

/* Material::Visibility, /root/reference/include/material/material.h (NormalVisible, VisibleNoShadows,
 * InvisibleShadowsOnly, Invisible) */
enum { YAFGPU_VIS_NORMAL = 0, YAFGPU_VIS_NO_SHADOWS = 1, YAFGPU_VIS_SHADOW_ONLY = 2, YAFGPU_VIS_INVISIBLE = 3 };

/* Flattened triangle soup.  Mirrors what TriangleObject holds per mesh
 * (/root/reference/include/object_geom/object_geom_mesh.h:84-144), concatenated over meshes:
 *   points   : vertex positions (3 floats each); when a mesh has orco the reference stores
 *              position and orco interleaved (getVertex(pa+1) is the orco, triangle.cc:47-49) and so does this array
 *   normals  : vertex normals after smoothing (3 floats each)
 *   uv       : uv_values_ (2 floats each)
 * Per triangle (n_tris entries, global order):
 *   tri_vidx : 3 indices into points (pa_, pb_, pc_, already offset to the global array)
 *   tri_nidx : 3 indices into normals, -1 = use geometric normal (na_, nb_, nc_; triangle.cc:38-40); may be NULL
 *   tri_uvidx: 3 indices into uv, may be NULL (no mesh has uv)
 *   tri_mat  : material index into yafgpu_scene_desc.materials
 *   tri_obj  : index into objects
 */
typedef struct yafgpu_object
{
	int32_t has_orco;      /* TriangleObject::has_orco_ */
	int32_t has_uv;        /* TriangleObject::has_uv_ (tri_uvidx valid for its triangles) */
	int32_t smooth;        /* is_smooth_ || normals_exported_ (triangle.cc:36) */
	int32_t light;         /* index of the light this object is tagged with (ObjectGeometric::setLight), -1 = none */
	int32_t first_tri;     /* first global triangle id */
	int32_t n_tris;
	int32_t obj_id;        /* reference ObjId_t, informational */
	int32_t pass_index;    /* obj_pass_index, informational */
} yafgpu_object;

typedef struct yafgpu_geometry
{
	const float *points;      size_t n_points;
	const float *normals;     size_t n_normals;
	const float *uv;          size_t n_uv;
	const int32_t *tri_vidx;
	const int32_t *tri_nidx;
	const int32_t *tri_uvidx;
	const int32_t *tri_mat;
	const int32_t *tri_obj;
	size_t n_tris;
	const yafgpu_object *objects; size_t n_objects;
} yafgpu_geometry;

/* ---- materials ------------------------------------------------------------------------- */

typedef enum yafgpu_material_type
{
	YAFGPU_MAT_SHINYDIFFUSE = 0,  /* "shinydiffusemat"  material_shiny_diffuse.cc:599-688 */
	YAFGPU_MAT_GLOSSY = 1,        /* "glossy"           material_glossy.cc:407-480 */
	YAFGPU_MAT_COATED_GLOSSY = 2, /* "coated_glossy"    material_coated_glossy.cc:465-545 */
	YAFGPU_MAT_GLASS = 3,         /* "glass"            material_glass.cc:311-379 */
	YAFGPU_MAT_ROUGH_GLASS = 4,   /* "rough_glass"      material_rough_glass.cc:322-401 */
	YAFGPU_MAT_BLEND = 5,         /* "blend_mat"        material_blend.cc:464-518 */
	YAFGPU_MAT_MIRROR = 6,        /* "mirror"           material_glass.cc (MirrorMaterial) */
	YAFGPU_MAT_NULL = 7,          /* "null"             material_glass.cc (NullMaterial) */
	YAFGPU_MAT_LIGHT = 8,         /* "light_mat"        material_simple.cc:38-73 */
	/* 9 and 10 are kernel classes internal to the library */
	YAFGPU_MAT_MASK = 11          /* "mask_mat"         material_mask.cc:30-198: u.blend.material1 / material2, blend_value = "threshold"; the "mask"
	                               * shader (YAFGPU_SLOT_BLEND) picks material2 where its scalar > threshold, else material1 */
} yafgpu_material_type;

/* ---- textures and shader nodes ------------------------------------- */

/* ImageTexture, /root/reference/src/texture/texture_image.cc:560-720 (factory), :75-230 (getColor, doMapping), :231-375 (interpolation).
 * Texels are handed over as ImageHandler::getPixel(x, y) returns them — linear RGB float after the handler's colour-space
 * decode and texture "optimization" storage round trip — so no image codec lives on this side of the ABI. */
typedef enum yafgpu_texture_type
{
	YAFGPU_TEX_IMAGE = 0,
	YAFGPU_TEX_BLEND = 1,    /* "blend"    texture_basic.cc:680-779: a gradient of the texture point (blend_type, use_flip_axis), optionally through a colour ramp.
	                          * Its factory sets no adjustments. */
	YAFGPU_TEX_RGB_CUBE = 2, /* "rgb_cube" texture_basic.cc:347-386: colour = the texture point clamped to [0,1]^3, with the adj_* adjustments */
	/* noise-based procedurals over NoiseGenerator + turbulence__ (noise_generator.cc:182-630,799-811); they need yafgpu_scene_desc.noise_tables */
	YAFGPU_TEX_CLOUDS = 3,   /* "clouds" texture_basic.cc:96-171 */
	YAFGPU_TEX_MARBLE = 4,   /* "marble" texture_basic.cc:177-250 */
	YAFGPU_TEX_WOOD = 5,     /* "wood"   texture_basic.cc:257-343 */
	YAFGPU_TEX_VORONOI = 6,  /* "voronoi" texture_basic.cc:391-503: weighted feature distances, seven metrics, cell colours */
	YAFGPU_TEX_MUSGRAVE = 7, /* "musgrave" texture_basic.cc:509-596 over the five fractals of noise_generator.cc:631-784 */
	YAFGPU_TEX_DISTORTED_NOISE = 8 /* "distorted_noise" texture_basic.cc:602-673 */
} yafgpu_texture_type;
/* VoronoiNoiseGenerator::DMetricType ("distance_metric": squared, manhattan, chebychev, minkovsky_half, minkovsky_four, minkovsky; else real) */
enum { YAFGPU_DIST_REAL = 0, YAFGPU_DIST_SQUARED = 1, YAFGPU_DIST_MANHATTAN = 2, YAFGPU_DIST_CHEBYCHEV = 3, YAFGPU_DIST_MINKOVSKY_HALF = 4,
       YAFGPU_DIST_MINKOVSKY_FOUR = 5, YAFGPU_DIST_MINKOVSKY = 6 };
/* "musgrave_type": multifractal, heteroterrain, hybridmf, ridgedmf; anything else = fBm */
enum { YAFGPU_MUSGRAVE_FBM = 0, YAFGPU_MUSGRAVE_MULTIFRACTAL = 1, YAFGPU_MUSGRAVE_HETEROTERRAIN = 2, YAFGPU_MUSGRAVE_HYBRIDMF = 3, YAFGPU_MUSGRAVE_RIDGEDMF = 4 };
/* what newNoise__(noise_type) constructs, texture_basic.cc:24-52: "blender", "stdperlin", "voronoi_f1|f2|f3|f4|f2f1|crackle" (DistReal metric), "cellnoise"; anything else = new Perlin */
enum { YAFGPU_NOISE_NEWPERLIN = 0, YAFGPU_NOISE_STDPERLIN = 1, YAFGPU_NOISE_BLENDER = 2, YAFGPU_NOISE_VORONOI_F1 = 3, YAFGPU_NOISE_VORONOI_F2 = 4, YAFGPU_NOISE_VORONOI_F3 = 5,
       YAFGPU_NOISE_VORONOI_F4 = 6, YAFGPU_NOISE_VORONOI_F2F1 = 7, YAFGPU_NOISE_VORONOI_CRACKLE = 8, YAFGPU_NOISE_CELL = 9 };
enum { YAFGPU_WAVE_SIN = 0, YAFGPU_WAVE_SAW = 1, YAFGPU_WAVE_TRI = 2 }; /* marble / wood "shape" */
#define YAFGPU_NOISE_TABLES_BYTES 13340
enum { YAFGPU_TEXBLEND_LINEAR = 0, YAFGPU_TEXBLEND_QUADRATIC = 1, YAFGPU_TEXBLEND_EASING = 2, YAFGPU_TEXBLEND_DIAGONAL = 3, YAFGPU_TEXBLEND_SPHERICAL = 4,
       YAFGPU_TEXBLEND_QUADRATIC_SPHERE = 5, YAFGPU_TEXBLEND_RADIAL = 6 }; /* BlendTexture::ProgressionType ("stype" lin/quad/ease/diag/sphere/halo|quad_sphere/radial) */
enum { YAFGPU_TEXINTERP_NONE = 0, YAFGPU_TEXINTERP_BILINEAR = 1, YAFGPU_TEXINTERP_BICUBIC = 2 }; /* "interpolate"; mipmap_* are refused */
enum { YAFGPU_TEXCLIP_EXTEND = 0, YAFGPU_TEXCLIP_CLIP = 1, YAFGPU_TEXCLIP_CLIPCUBE = 2, YAFGPU_TEXCLIP_REPEAT = 3, YAFGPU_TEXCLIP_CHECKER = 4 }; /* ImageTexture::TexClipMode */
enum { YAFGPU_CS_RAW_MANUAL_GAMMA = 1, YAFGPU_CS_LINEAR_RGB = 2, YAFGPU_CS_SRGB = 3, YAFGPU_CS_XYZ_D65 = 4 }; /* ColorSpace, include/common/color.h:33-39 */

typedef struct yafgpu_texture
{
	int32_t type;            /* yafgpu_texture_type */
	int32_t width, height;   /* ImageHandler::getWidth/getHeight */
	const float *rgba;       /* width*height*4 floats: getPixel(x, y) at rgba + 4 * (y * width + x) */
	int32_t interpolate;     /* YAFGPU_TEXINTERP_* */
	int32_t clip_mode;       /* "clipping" (YAFGPU_TEXCLIP_*; the factory's fallback for unknown strings is repeat) */
	int32_t xrepeat, yrepeat;/* "xrepeat", "yrepeat" */
	int32_t mirror_x, mirror_y; /* "mirror_x", "mirror_y" */
	int32_t rot90;           /* "rot90" */
	float cropmin_x, cropmin_y, cropmax_x, cropmax_y; /* "cropmin_x" ... (narrowed to float by setCrop) */
	int32_t checker_even, checker_odd; /* "even_tiles", "odd_tiles" */
	float checker_dist;      /* "checker_dist" */
	int32_t normalmap;       /* "normalmap" */
	int32_t color_space;     /* YAFGPU_CS_* the handler decoded from (HDR/EXR: linear); used by getRawColor to re-encode */
	float gamma;             /* "gamma" */
	float adj_intensity, adj_contrast, adj_saturation; /* "adj_intensity", "adj_contrast", "adj_saturation" (image, rgb_cube) */
	float adj_hue;           /* "adj_hue" / 60.f in float, i.e. Texture::adj_hue_ as setAdjustments stores it (texture.h:154) */
	float adj_mult_red, adj_mult_green, adj_mult_blue;
	int32_t adj_clamp;       /* "adj_clamp" */
	int32_t blend_type;      /* blend: YAFGPU_TEXBLEND_* */
	int32_t use_flip_axis;   /* blend: "use_flip_axis" */
	/* clouds / marble / wood: the members of the constructed texture object */
	int32_t noise_type;      /* YAFGPU_NOISE_* */
	int32_t noise_depth;     /* "depth" (octaves - 1 of turbulence__) */
	float noise_size;        /* "size" */
	int32_t noise_hard;      /* "hard" */
	float color1[3], color2[3]; /* "color1", "color2" */
	int32_t clouds_bias;     /* clouds "bias": 0 none, 1 positive, 2 negative */
	float turbulence;        /* marble, wood "turbulence" */
	float sharpness;         /* marble: sharpness_ as the constructor stores it (1, or 1 / "sharpness" when that is > 1) */
	int32_t wave_shape;      /* marble, wood: YAFGPU_WAVE_* */
	int32_t wood_rings;      /* wood: "wood_type" == "rings" */
	/* voronoi (noise_size = "size"; color1/color2 are read by the factory and never used) */
	float vor_weight[4];     /* "weight1" .. "weight4" */
	float mk_exponent;       /* "mk_exponent" */
	float intensity_scale;   /* iscale_ of the constructed texture: musgrave = "intensity"; voronoi = "intensity" / (|w1|+|w2|+|w3|+|w4|), float, summed left to right (0 when all weights are 0) */
	int32_t distance_metric; /* YAFGPU_DIST_* */
	int32_t color_type;      /* the factory's "Rgbype" string (sic): 0 intensity, 1 col1, 2 col2, 3 col3 */
	/* musgrave (noise_type, noise_size, color1/2 as above) */
	int32_t musgrave_type;   /* YAFGPU_MUSGRAVE_* */
	float mus_h, mus_lacunarity, mus_octaves, mus_offset, mus_gain; /* "H", "lacunarity", "octaves", "offset", "gain" */
	/* distorted_noise: noise_type = "noise_type1" (the distortion), noise_type2 = "noise_type2" (the distorted basis) */
	float distort;           /* "distort" */
	int32_t noise_type2;
	/* colour ramp of a procedural texture ("use_color_ramp" + "ramp_*", textureReadColorRamp__ texture_basic.cc:54-90; ColorRamp color_ramp.cc:31-164): the texture's
	 * scalar picks a colour between the two items around it.  ramp_items = 0: no ramp.  The items are r, g, b, a, position, sorted by position as ColorRamp::addItem
	 * keeps them; the first position must be <= 0 and the last >= 1 (getColorInterpolated reads in front of / behind the array otherwise). */
	int32_t ramp_items;
	int32_t ramp_mode;          /* YAFGPU_RAMP_RGB / HSV / HSL ("ramp_color_mode") */
	int32_t ramp_interpolation; /* 0 constant, 1 linear ("ramp_interpolation": CONSTANT, anything else is linear) */
	int32_t ramp_hue_interpolation; /* YAFGPU_RAMPHUE_* ("ramp_hue_interpolation": NEAR, FAR, CW, CCW) */
	const float *ramp;          /* ramp_items * 5 floats */
} yafgpu_texture;
enum { YAFGPU_RAMP_RGB = 0, YAFGPU_RAMP_HSV = 1, YAFGPU_RAMP_HSL = 2 };
enum { YAFGPU_RAMPHUE_NEAR = 0, YAFGPU_RAMPHUE_FAR = 1, YAFGPU_RAMPHUE_CW = 2, YAFGPU_RAMPHUE_CCW = 3 };

typedef enum yafgpu_node_type
{
	YAFGPU_NODE_TEXTURE_MAPPER = 0, /* "texture_mapper" shader_node_basic.cc:28-420 */
	YAFGPU_NODE_VALUE = 1,          /* "value"          shader_node_basic.cc:426-446 */
	YAFGPU_NODE_MIX = 2,            /* "mix"            shader_node_basic.cc:452-705 (mode picks Mix/Add/Mult/Sub/Screen/Diff/Dark/Light/Overlay) */
	YAFGPU_NODE_LAYER = 3           /* "layer"          shader_node_layer.cc:25-244 */
} yafgpu_node_type;
enum { YAFGPU_TEXCO_UV = 0, YAFGPU_TEXCO_GLOBAL = 1, YAFGPU_TEXCO_ORCO = 2, YAFGPU_TEXCO_TRANSFORMED = 3, YAFGPU_TEXCO_NORMAL = 4,
       YAFGPU_TEXCO_REFLECT = 5, YAFGPU_TEXCO_WINDOW = 6, YAFGPU_TEXCO_STICK = 7, YAFGPU_TEXCO_STRESS = 8, YAFGPU_TEXCO_TANGENT = 9 }; /* TextureMapperNode::Coords */
enum { YAFGPU_TEXPROJ_PLAIN = 0, YAFGPU_TEXPROJ_CUBE = 1, YAFGPU_TEXPROJ_TUBE = 2, YAFGPU_TEXPROJ_SPHERE = 3 };                   /* TextureMapperNode::Projection */
/* MixModes, include/shader/shader_node.h:66 */
enum { YAFGPU_MIX_MIX = 0, YAFGPU_MIX_ADD, YAFGPU_MIX_MULT, YAFGPU_MIX_SUB, YAFGPU_MIX_SCREEN, YAFGPU_MIX_DIV, YAFGPU_MIX_DIFF, YAFGPU_MIX_DARK, YAFGPU_MIX_LIGHT, YAFGPU_MIX_OVERLAY };

/* One <list_element element="shader_node">.  Fields are the values the node factories / configInputs read (after defaults). */
typedef struct yafgpu_shader_node
{
	int32_t type;            /* yafgpu_node_type */
	/* inputs: index into the owning material's node list, -1 = the parameter is absent.
	 *   mix:   input1 = "input1", input2 = "input2", factor = "factor"
	 *   layer: input1 = "input",  input2 = "upper_layer" */
	int32_t input1, input2, factor;
	int32_t mode;            /* mix, layer: "mode" */
	/* texture_mapper (factory shader_node_basic.cc:344-420) */
	int32_t texture;         /* index into yafgpu_scene_desc.textures ("texture") */
	int32_t texco;           /* "texco"   (YAFGPU_TEXCO_*, default global) */
	int32_t mapping;         /* "mapping" (YAFGPU_TEXPROJ_*, default plain; only honoured for discrete textures) */
	int32_t proj_x, proj_y, proj_z; /* "proj_x/y/z" (clamped to 0..3 by the factory) */
	float scale[3];          /* "scale" */
	float offset[3];         /* "offset" (as given; the factory doubles it) */
	float bump_strength;     /* "bump_strength" */
	int32_t do_scalar;       /* "do_scalar" */
	float transform[16];     /* "transform", row-major Matrix4 (identity by default) */
	/* value: color = "color" + "alpha", scalar = "scalar".  mix: color = "color1", color2 = "color2", scalar = "value"
	 * (constant inputs of a mix node have no scalar part in the reference — val_1_/val_2_ are never set — it is taken as 0).
	 * layer: color = "def_col" (alpha 1), color2 = "upper_color", scalar = "def_val", scalar2 = "upper_value" */
	float color[4], color2[4];
	float scalar, scalar2;
	/* layer (factory shader_node_layer.cc:203-244) */
	float colfac, valfac;    /* "colfac", "valfac" */
	int32_t do_color, layer_do_scalar, color_input, use_alpha; /* "do_color", "do_scalar", "color_input", "use_alpha" */
	int32_t no_rgb, stencil, negative;                         /* "noRGB", "stencil", "negative" */
} yafgpu_shader_node;

/* "<slot>_shader" material parameters (e.g. material_shiny_diffuse.cc:696-724, material_glossy.cc, material_coated_glossy.cc,
 * material_glass.cc, material_rough_glass.cc, material_blend.cc) */
enum { YAFGPU_SLOT_DIFFUSE = 0,       /* "diffuse_shader" */
       YAFGPU_SLOT_MIRROR_COLOR = 1,  /* "mirror_color_shader" */
       YAFGPU_SLOT_BUMP = 2,          /* "bump_shader" */
       YAFGPU_SLOT_MIRROR = 3,        /* "mirror_shader" */
       YAFGPU_SLOT_TRANSPARENCY = 4,  /* "transparency_shader" */
       YAFGPU_SLOT_TRANSLUCENCY = 5,  /* "translucency_shader" */
       YAFGPU_SLOT_SIGMA_OREN = 6,    /* "sigma_oren_shader" */
       YAFGPU_SLOT_DIFFUSE_REFL = 7,  /* "diffuse_refl_shader" */
       YAFGPU_SLOT_IOR = 8,           /* "IOR_shader" */
       YAFGPU_SLOT_WIREFRAME = 9,     /* "wireframe_shader" */
       YAFGPU_SLOT_GLOSSY = 10,       /* "glossy_shader" */
       YAFGPU_SLOT_GLOSSY_REFLECT = 11, /* "glossy_reflect_shader" */
       YAFGPU_SLOT_EXPONENT = 12,     /* "exponent_shader" */
       YAFGPU_SLOT_FILTER_COLOR = 13, /* "filter_color_shader" */
       YAFGPU_SLOT_ROUGHNESS = 14,    /* "roughness_shader" */
       YAFGPU_SLOT_BLEND = 15,        /* blend_mat / mask_mat "mask" */
       YAFGPU_N_SLOTS = 16 };

typedef struct yafgpu_mat_shinydiffuse
{
	float color[3];          /* "color" */
	float mirror_color[3];   /* "mirror_color" */
	float diffuse_reflect;   /* "diffuse_reflect" */
	float specular_reflect;  /* "specular_reflect" (mirror strength) */
	float transparency;      /* "transparency" */
	float translucency;      /* "translucency" */
	float emit;              /* "emit" */
	float ior;               /* "IOR" (used only with fresnel_effect) */
	int32_t fresnel_effect;  /* "fresnel_effect" */
	float transmit_filter;   /* "transmit_filter" */
	int32_t oren_nayar;      /* "diffuse_brdf" == "oren_nayar" */
	float sigma;             /* "sigma" */
} yafgpu_mat_shinydiffuse;

typedef struct yafgpu_mat_glossy /* also the base of coated_glossy */
{
	float color[3];          /* "color" (glossy colour) */
	float diffuse_color[3];  /* "diffuse_color" */
	float mirror_color[3];   /* coated only: "mirror_color" */
	float diffuse_reflect;   /* "diffuse_reflect" */
	float glossy_reflect;    /* "glossy_reflect" */
	float specular_reflect;  /* coated only: "specular_reflect" */
	float exponent;          /* "exponent" */
	float exp_u, exp_v;      /* "exp_u", "exp_v" */
	float ior;               /* coated only: "IOR" */
	int32_t as_diffuse;      /* "as_diffuse" */
	int32_t anisotropic;     /* "anisotropic" */
	int32_t oren_nayar;      /* "diffuse_brdf" == "Oren-Nayar" */
	float sigma;             /* "sigma" */
} yafgpu_mat_glossy;

typedef struct yafgpu_mat_glass /* glass and rough_glass */
{
	float ior;               /* "IOR" */
	float filter_color[3];   /* "filter_color" */
	float transmit_filter;   /* "transmit_filter" */
	float mirror_color[3];   /* "mirror_color" */
	float dispersion_power;  /* "dispersion_power" */
	int32_t fake_shadows;    /* "fake_shadows" */
	float alpha;             /* rough_glass only: "alpha" */
	int32_t has_absorption;  /* the factory created the Beer volume handler (Material::vol_i_): "absorption" given and < 1 in some channel AND the
	                          * material's ParamMap carries a "name" parameter (material_glass.cc:376-399); without it "absorption" has no effect */
	float absorption[3];     /* "absorption" */
	float absorption_dist;   /* "absorption_dist" (1 when absent) */
} yafgpu_mat_glass;

typedef struct yafgpu_mat_blend
{
	int32_t material1;       /* index of "material1" */
	int32_t material2;       /* index of "material2" */
	float blend_value;       /* blend_mat "blend_value"; mask_mat "threshold" */
} yafgpu_mat_blend;

typedef struct yafgpu_mat_simple /* mirror: color*reflect; light_mat: color*power, double_sided */
{
	float color[3];
	float strength;          /* mirror "reflect" / light_mat "power" */
	int32_t double_sided;
} yafgpu_mat_simple;

typedef struct yafgpu_material
{
	int32_t type;                 /* yafgpu_material_type */
	int32_t visibility;           /* "visibility" */
	int32_t receive_shadows;      /* "receive_shadows" */
	int32_t flat_material;        /* "flat_material" (shinydiffuse) */
	int32_t additional_depth;     /* "additionaldepth" */
	float transparent_bias_factor;            /* "transparentbias_factor" */
	int32_t transparent_bias_multiply_raydepth; /* "transparentbias_multiply_raydepth" */
	float sampling_factor;        /* "samplingfactor" */
	int32_t mat_pass_index;       /* "mat_pass_index" */
	/* shader nodes of this material (NodeMaterial::loadNodes material_node.cc:140-214): nodes[node_first .. node_first+node_count)
	 * of yafgpu_scene_desc.nodes, in DECLARATION order (the <list_element> order); the library derives the evaluation order
	 * (solveNodesOrder material_node.cc:85-103) and the colour / bump node lists the material factories build.  node_count = 0:
	 * no nodes (also what the reference is left with when loadNodes fails, e.g. on a texture that could not be loaded). */
	int32_t node_first, node_count;
	int32_t reserved;
	/* "<slot>_shader" parameters: index into this material's node list, -1 = not given (YAFGPU_SLOT_*) */
	int32_t shader[YAFGPU_N_SLOTS];
	union
	{
		yafgpu_mat_shinydiffuse shinydiffuse;
		yafgpu_mat_glossy glossy;
		yafgpu_mat_glass glass;
		yafgpu_mat_blend blend;
		yafgpu_mat_simple simple;
		float pad_[32];
	} u;
} yafgpu_material;

/* ---- lights ---------------------------------------------------------------------------- */

typedef enum yafgpu_light_type
{
	YAFGPU_LIGHT_POINT = 0,      /* "pointlight"  light_point.cc:101-133 */
	YAFGPU_LIGHT_SUN = 1,        /* "sunlight"    light_sun.cc:95-125 */
	YAFGPU_LIGHT_AREA = 2,       /* "arealight"   light_area.cc:174-213 */
	YAFGPU_LIGHT_MESH = 3,       /* "meshlight"   light_meshlight.cc:228-263 */
	YAFGPU_LIGHT_BACKGROUND = 4, /* "bglight"     light_background.cc:251-283 (created by a background with ibl=true) */
	YAFGPU_LIGHT_DIRECTIONAL = 5,/* "directional" light_directional.cc:112-158 */
	YAFGPU_LIGHT_SPOT = 6,       /* "spotlight"   light_spot.cc:251-295 */
	YAFGPU_LIGHT_SPHERE = 7,     /* "spherelight" light_sphere.cc:163-198 (without a linked sphere primitive: "object" must be 0) */
	YAFGPU_LIGHT_IES = 9,        /* "ieslight"    light_ies.cc:184-233: a point source whose intensity follows the photometric web of an IESNA LM-63 file;
	                              * the caller passes the FILE CONTENTS (ies_data / ies_size), the library parses them as IesData::parseIesFile does
	                              * (include/utility/util_ies.h:131-379).  A file the reference's parser rejects makes its factory return no light at
	                              * all: such a light must not be passed (upload fails with YAFGPU_ERR_INVALID) */
	YAFGPU_LIGHT_BGPORTAL = 8    /* "bgPortalLight" light_background_portal.cc:236-266: the mesh "object" is a window that radiates the background
	                              * (colour = Background::eval(direction) * power); the mesh must be present in the geometry — the reference's
	                              * kd-tree is built before the light's init() hides it (scene.cc:788-831 precede the light loop) */
} yafgpu_light_type;

typedef struct yafgpu_light
{
	int32_t type;            /* yafgpu_light_type */
	int32_t light_enabled;   /* "light_enabled" */
	int32_t cast_shadows;    /* "cast_shadows" */
	int32_t photon_only;     /* "photon_only" (such lights never illuminate; light_area.cc:69) */
	int32_t samples;         /* "samples" (sun/area/mesh/bg) */
	int32_t object;          /* area/mesh: index into yafgpu_geometry.objects of the "object" param, -1 = none */
	int32_t double_sided;    /* mesh: "double_sided" */
	int32_t abs_intersect;   /* bg: "abs_intersect" */
	float color[3];          /* "color" */
	float power;             /* "power" */
	float from[3];           /* point/spot/sphere/ies/finite directional: "from"; sun: "direction"; area: "corner" */
	float point1[3];         /* area: "point1"; spot, ies: "to"; directional: "direction" */
	float point2[3];         /* area: "point2" */
	float angle;             /* sun: "angle" (degrees) */
	float clamp_intersect;   /* bg: ibl_clamp_sampling */
	int32_t shoot_caustic;   /* "with_caustic" (bg light: gates caustic-path background hits) */
	int32_t shoot_diffuse;   /* "with_diffuse" */
	float radius;            /* sphere: "radius"; finite directional: "radius" */
	float cone_angle;        /* spot: "cone_angle" (degrees) */
	float blend;             /* spot: "blend" */
	float shadow_fuzzy;      /* spot: "shadowFuzzyness" */
	int32_t soft_shadows;    /* spot, ies: "soft_shadows" (then it is an area-like light with "samples") */
	int32_t infinite;        /* directional: "infinite" */
	const char *ies_data;    /* ies: contents of the file named by "file" (need not be NUL terminated) */
	size_t ies_size;
} yafgpu_light;

/* ---- camera ---------------------------------------------------------------------------- */

typedef enum yafgpu_camera_type
{
	YAFGPU_CAM_PERSPECTIVE = 0,     /* "perspective"     camera_perspective.cc */
	YAFGPU_CAM_ARCHITECT = 1,       /* "architect"       camera_architect.cc (perspective with a vertical image plane) */
	YAFGPU_CAM_ORTHOGRAPHIC = 2,    /* "orthographic"    camera_orthographic.cc */
	YAFGPU_CAM_ANGULAR = 3,         /* "angular"         camera_angular.cc */
	YAFGPU_CAM_EQUIRECTANGULAR = 4  /* "equirectangular" camera_equirectangular.cc */
} yafgpu_camera_type;
enum { YAFGPU_AP_EQUIDISTANT = 0, YAFGPU_AP_ORTHOGRAPHIC = 1, YAFGPU_AP_STEREOGRAPHIC = 2, YAFGPU_AP_EQUISOLID = 3, YAFGPU_AP_RECTILINEAR = 4 };
enum { YAFGPU_BK_DISK1 = 0, YAFGPU_BK_DISK2 = 1, YAFGPU_BK_TRI = 3, YAFGPU_BK_SQR = 4, YAFGPU_BK_PENTA = 5, YAFGPU_BK_HEXA = 6, YAFGPU_BK_RING = 7 };
enum { YAFGPU_BB_NONE = 0, YAFGPU_BB_CENTER = 1, YAFGPU_BB_EDGE = 2 };

/* "perspective": camera_perspective.cc:206-252 and Camera::Camera camera.cc:46-66 */
typedef struct yafgpu_camera
{
	int32_t type;
	int32_t resx, resy;      /* "resx", "resy" */
	float from[3], to[3], up[3];
	float focal;             /* "focal" */
	float aperture;          /* "aperture" */
	float dof_distance;      /* "dof_distance" */
	float aspect_ratio;      /* "aspect_ratio" */
	float near_clip;         /* "nearClip" */
	float far_clip;          /* "farClip" */
	int32_t bokeh_type;      /* "bokeh_type" */
	int32_t bokeh_bias;      /* "bokeh_bias" */
	float bokeh_rotation;    /* "bokeh_rotation" */
	float scale;             /* orthographic: "scale" */
	double angle;            /* angular: "angle" (degrees; the factory converts to radians in double, camera_angular.cc:126) */
	double max_angle;        /* angular: "max_angle" (degrees; the factory defaults it to "angle") */
	int32_t circular;        /* angular: "circular" */
	int32_t mirrored;        /* angular: "mirrored" */
	int32_t projection;      /* angular: "projection" (YAFGPU_AP_*) */
} yafgpu_camera;

/* ---- background ------------------------------------------------------------------------ */

typedef enum yafgpu_background_type
{
	YAFGPU_BG_NONE = 0,
	YAFGPU_BG_CONSTANT = 1,   /* "constant"    background_constant.cc:52-86 */
	YAFGPU_BG_TEXTURE = 2,    /* "textureback" background_texture.cc:90-165 (image given as decoded linear float RGB) */
	YAFGPU_BG_GRADIENT = 3    /* "gradientback" background_gradient.cc:66-117 */
} yafgpu_background_type;

typedef struct yafgpu_background
{
	int32_t type;
	float color[3];          /* constant: "color" (before *power) */
	float power;             /* "power" */
	int32_t ibl;             /* "ibl": a yafgpu_light of type BACKGROUND must then be present in lights[] */
	int32_t with_caustic;    /* "with_caustic" -> Background::shootsCaustic() */
	int32_t mapping_angular; /* texture: "mapping" == probe/angular */
	float rotation;          /* texture: "rotation" (degrees) */
	/* texture: pixels as ImageHandler::getPixel returns them, row-major, y down, RGB float32 */
	const float *tex_rgb; int32_t tex_w, tex_h;
	int32_t tex_interp;      /* 0 = none, 1 = bilinear (texture "interpolate") */
	/* gradient: "horizon_color", "zenith_color", "horizon_ground_color", "zenith_ground_color" (before *power; the factory
	 * defaults the ground colours to the sky colours) */
	float horizon_color[3], zenith_color[3], horizon_ground_color[3], zenith_ground_color[3];
} yafgpu_background;

/* ---- film, AA and integrator settings -------------------------------------------------- */

enum { YAFGPU_FILTER_BOX = 0, YAFGPU_FILTER_MITCHELL = 1, YAFGPU_FILTER_GAUSS = 2, YAFGPU_FILTER_LANCZOS = 3 };

/* RenderEnvironment::createImageFilm environment.cc:456-584 and setupScene :679-813 */
typedef struct yafgpu_film_desc
{
	int32_t width, height;   /* "width", "height" */
	int32_t xstart, ystart;  /* "xstart", "ystart" */
	float aa_pixelwidth;     /* "AA_pixelwidth" */
	int32_t filter_type;     /* "filter_type" */
	int32_t tile_size;       /* "tile_size" (multi-GPU interleave unit) */
	int32_t premult;         /* "premult" */
	float aa_clamp_samples;  /* "AA_clamp_samples" */
	float aa_clamp_indirect; /* "AA_clamp_indirect" */
	int32_t aa_passes;       /* "AA_passes": only whether >1 matters to the sampler (integrator_tiled.cc:394) */
	float aa_light_sample_multiplier;    /* current value of TiledIntegrator::aa_light_sample_multiplier_ */
	float aa_indirect_sample_multiplier; /* current value of aa_indirect_sample_multiplier_ */
	uint32_t base_sampling_offset;       /* ImageFilm::getBaseSamplingOffset() (imagefilm.h:127) */
} yafgpu_film_desc;

typedef enum yafgpu_integrator_type
{
	YAFGPU_INT_DIRECTLIGHT = 0, /* "directlighting" integrator_direct_light.cc:196-250 */
	YAFGPU_INT_PATHTRACING = 1  /* "pathtracing"    integrator_path_tracer.cc:349-423 */
} yafgpu_integrator_type;

typedef struct yafgpu_integrator_desc
{
	int32_t type;
	int32_t raydepth;        /* "raydepth" */
	int32_t transp_shad;     /* "transpShad" */
	int32_t shadow_depth;    /* "shadowDepth" */
	int32_t path_samples;    /* "path_samples" */
	int32_t bounces;         /* "bounces" */
	int32_t rr_min_bounces;  /* "russian_roulette_min_bounces" */
	int32_t no_recursive;    /* "no_recursive" */
	int32_t trace_caustics;  /* caustic_type in {path, both} (integrator_path_tracer.cc:87); photon maps are refused */
	int32_t bg_transp;       /* "bg_transp" */
	int32_t bg_transp_refract; /* "bg_transp_refract": alpha = Material::getAlpha folded along the refraction chain, no background colour (integrator_path_tracer.cc:319-328) */
	int32_t do_ao;           /* "do_AO" */
	int32_t ao_samples;      /* "AO_samples" */
	float ao_distance;       /* "AO_distance" */
	float ao_color[3];       /* "AO_color" */
	/* Scene::shadow_bias_, ray_min_dist_ and their auto flags (scene.h:230-234, scene.cc:823-824) */
	float shadow_bias; int32_t shadow_bias_auto;
	float ray_min_dist; int32_t ray_min_dist_auto;
} yafgpu_integrator_desc;

typedef struct yafgpu_scene_desc
{
	uint32_t abi_version;    /* YAFGPU_ABI_VERSION */
	yafgpu_geometry geometry;
	const yafgpu_material *materials; size_t n_materials;
	const yafgpu_light *lights;       size_t n_lights;   /* in Scene::lights_ order */
	const yafgpu_texture *textures;   size_t n_textures; /* referenced by texture_mapper nodes */
	const yafgpu_shader_node *nodes;  size_t n_nodes;    /* all materials' nodes; yafgpu_material.node_first/node_count slice it */
	yafgpu_camera camera;
	yafgpu_background background;
	yafgpu_film_desc film;
	yafgpu_integrator_desc integrator;
	/* CIE 1931 colour matching functions, 471 rows (360..830 nm, 1 nm) of x, y, z — the table behind wl2RgbFromCie__ (spectrum.cc:276-289).
	 * Needed only when a glass / rough_glass material has dispersion_power > 0 (upload fails without it); the reference-side adaptor fills it
	 * from the reference's own wl2Xyz__, the Python mirror from libyafaray_b200/data/cie_xyz_360_830.npy (tools/make_cie_table.py). */
	const float *cie_xyz; size_t n_cie_rows;
	/* Lattice tables of the reference's noise bases, required when a clouds / marble / wood texture is present: YAFGPU_NOISE_TABLES_BYTES bytes laid out as
	 * f32 hashpntf[768], f32 hashvectf[768], f32 stdp_g[514][3], u8 hash[512], u8 stdp_p[514], 2 pad bytes (noise_generator.cc:25-320; they are file-static
	 * literal data there).  The adaptor compiles them in from a header that oracle/ref_driver/noise_dump.cc generates out of the reference's own source at
	 * build time; the Python mirror reads libyafaray_b200/data/noise_tables.bin, written by the same program (tools/make_noise_tables.py). */
	const void *noise_tables; size_t noise_tables_bytes;
} yafgpu_scene_desc;

/* ---- rays and hits (parity probe + traversal microbench) -------------------------------- */

typedef struct yafgpu_ray { float from[3]; float tmin; float dir[3]; float tmax; } yafgpu_ray; /* tmax < 0 = infinite (scene.cc:900) */
typedef struct yafgpu_hit { float t, u, v; int32_t tri; } yafgpu_hit;                          /* tri = global triangle id, -1 = miss; u,v = Möller–Trumbore b_1,b_2 */

enum { YAFGPU_TRACE_CLOSEST = 0,      /* Scene::intersect          scene.cc:896-927 */
       YAFGPU_TRACE_SHADOW = 1,       /* Scene::isShadowed opaque  scene.cc:962-995  (hit.tri = -1 unshadowed, else the blocker found; t = 1) */
       YAFGPU_TRACE_SHADOW_TRANSP = 2 /* Scene::isShadowed transp. scene.cc:997-1035 (u,v,t reused: filt r,g,b in t,u,v) */ };

/* ---- render pass ----------------------------------------------------------------------- */

/* One call of TiledIntegrator::renderPass (integrator_tiled.cc:261-307) for this context's share of the image. */
typedef struct yafgpu_pass_desc
{
	int32_t n_samples;       /* samples per pixel in this pass */
	int32_t offset;          /* pass_offs: sampling offset incl. base offset (integrator_tiled.cc:263,335) */
	int32_t adaptive;        /* 1: only pixels flagged in resample_mask are sampled (imagefilm.cc:917-920); 2: the mask is the one the last yafgpu_next_pass computed */
	int32_t aa_pass_number;
	const uint8_t *resample_mask; /* adaptive only: width*height bytes (host), non-zero = sample this pixel */
	/* sub-range of the pass's samples to render now: sample indices [sample_begin, sample_end); both 0 = all.
	 * Lets a caller split one pass into bounded steps without changing any sample's random numbers. */
	int32_t sample_begin, sample_end;
	/* image partition for multi-GPU: this context renders the blocks (tx, ty) with (tx + ty * stride) % part_count == part_index of a raster grid of
	 * min(tile_size, 8)-pixel square blocks (the shape of ImageSplitter's tile grid, imagesplitter.cc:37-50); stride = the integer nearest 0.618 * part_count that is
	 * coprime with it, so that every part walks through all columns.  The parts of one pass are disjoint and cover the image.  part_count <= 1 = whole image. */
	int32_t part_index, part_count;
} yafgpu_pass_desc;

/* ImageFilm::setAaNoiseParams / setAaThreshold (imagefilm.h), i.e. the <render> parameters AA_threshold, AA_dark_detection_type (none|linear|curve),
 * AA_dark_threshold_factor, AA_variance_edge_size, AA_variance_pixels, AA_detect_color_noise */
enum { YAFGPU_DARK_NONE = 0, YAFGPU_DARK_LINEAR = 1, YAFGPU_DARK_CURVE = 2 };
typedef struct yafgpu_aa_noise_params
{
	float threshold;
	int32_t dark_detection_type;
	float dark_threshold_factor;
	int32_t variance_edge_size, variance_pixels;
	int32_t detect_color_noise;
} yafgpu_aa_noise_params;

typedef struct yafgpu_stats
{
	uint64_t rays_camera, rays_extension, rays_shadow, rays_shadow_transp, rays_light_probe;
	uint64_t camera_samples;
	uint64_t kernel_launches;
	uint64_t bvh_nodes, bvh_bytes, tri_bytes;
	uint64_t nodes_visited, tris_tested;   /* filled only when the context was created with YAFGPU_FLAG_COUNT_TRAVERSAL */
	double ms_build, ms_trace, ms_shade, ms_film, ms_total; /* CUDA-event times accumulated since the last reset */
	uint64_t waves, wave_retries;
	/* filled only with YAFGPU_FLAG_TIME_STAGES: per-kernel-class CUDA-event time and NON-EMPTY launch counts
	 * (ms_trace = closest-hit kernel, ms_shadow = any-hit kernel, ms_shade = shading kernel, ms_film = raygen + film splat) */
	double ms_shadow;
	uint64_t launches_trace, launches_shadow, launches_shade;
	uint64_t rays_in_timed_trace, rays_in_timed_shadow;
	/* queue lengths of the most recent wave, per wavefront iteration (first 16): tasks traced+shaded, shadow rays traced */
	uint32_t last_wave_tasks[16], last_wave_shadow[16];
	/* accelerator shape: bytes per node (128: 4-wide fp32 boxes), children per node, levels of the wide tree */
	uint32_t bvh_node_bytes, bvh_width, bvh_levels, reserved0;
	/* COUNT_TRAVERSAL: nodes_visited / tris_tested above cover closest-hit rays, these the shadow (any-hit) rays */
	uint64_t nodes_visited_shadow, tris_tested_shadow;
	/* yafgpu_group_get_stats only: film reductions queued so far and the duration of the most recent one that was read back */
	uint64_t film_reduces; double ms_film_reduce;
	/* camera samples whose radiance came back NaN/Inf and were left out of the film (see option "keep_nonfinite_samples") */
	uint64_t nonfinite_samples;
} yafgpu_stats;

enum { YAFGPU_FLAG_COUNT_TRAVERSAL = 1, YAFGPU_FLAG_TIME_STAGES = 2,
       YAFGPU_FLAG_SHARE_DEVICE = 4 /* yafgpu_group_create only: the device list may name a GPU more than once (several members share it; exercises the
                                       partition + film-sum path on a single-GPU box) */ };

/* ---- entry points ---------------------------------------------------------------------- */

/* Create a context on CUDA device `device`.  Fails (no fallback) when the device is not usable. */
int yafgpu_create(yafgpu_ctx **out, int device, unsigned flags);
void yafgpu_destroy(yafgpu_ctx *ctx);
const char *yafgpu_last_error(const yafgpu_ctx *ctx); /* ctx may be NULL: error of the last failed yafgpu_create */
int yafgpu_abi_version(void);

/* Replaces the geometry branch of Scene::update() (scene.cc:788-831): copies the scene to HBM, builds the BVH
 * (instead of `new TriKdTree`, kdtree_triangle.cc:76-157), initialises lights (Light::init) and the film (ImageFilm::init). */
int yafgpu_upload_scene(yafgpu_ctx *ctx, const yafgpu_scene_desc *desc);
/* Scene::getSceneBound() (scene.cc): 6 floats min xyz, max xyz */
int yafgpu_scene_bound(yafgpu_ctx *ctx, float *bound6);

/* Scene::intersect / Scene::isShadowed on a batch of host rays (H2D, trace, D2H).  shadow_depth = max_depth of scene.cc:997. */
int yafgpu_trace(yafgpu_ctx *ctx, const yafgpu_ray *rays, size_t n, yafgpu_hit *hits, int mode, int shadow_depth);
/* Same with device pointers (no copies): for microbenchmarks that keep rays resident in HBM. */
int yafgpu_trace_device(yafgpu_ctx *ctx, const void *d_rays, size_t n, void *d_hits, int mode, int shadow_depth);
/* yafgpu_trace for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.  Same results as yafgpu_trace
 * for YAFGPU_TRACE_CLOSEST / YAFGPU_TRACE_SHADOW, but the device copies of `rays` and `hits` and the kernel's work counter live in
 * a file-scope device array of the library (3 MiB + 256 B per GPU that loads it, held until the process exits) instead of the
 * memory pool: the call performs no cudaMalloc / cudaFree / cudaMallocAsync and calls no allocating library routine.
 * At most YAFGPU_TRACE_STATIC_MAX_RAYS rays per call (more: YAFGPU_ERR_INVALID, nothing is traced).  All calls on one GPU share the
 * array, which keeps the last call's contents; callers inside one process are serialised by the library, but work that overlaps on
 * the GPU from elsewhere (another host process driving the same array is impossible, each has its own copy; a caller that bypasses this
 * entry point is not) would race.  Synchronous like yafgpu_trace: returns with `hits` filled. */
#define YAFGPU_TRACE_STATIC_MAX_RAYS 65536
int yafgpu_trace_static(yafgpu_ctx *ctx, const yafgpu_ray *rays, size_t n, yafgpu_hit *hits, int mode);

/* ImageFilm::init (imagefilm.cc:219-268): zero the film accumulators. */
int yafgpu_film_clear(yafgpu_ctx *ctx);
/* TiledIntegrator::renderPass: trace + shade + ImageFilm::addSample for the pass (or the sample sub-range given).
 * Returns when the pass is in the film accumulators (the wave scheduler reads queue counters back after every wave). */
int yafgpu_render_pass(yafgpu_ctx *ctx, const yafgpu_pass_desc *pass);
/* Wait for all queued work; returns the first asynchronous error. */
/* ImageFilm::nextPass's resampling decision (imagefilm.cc:306-424) evaluated ON THE DEVICE from the film accumulators: flags every pixel whose normalised colour
 * differs from a right / lower / diagonal neighbour by at least the (dark-scaled) threshold, the variance windows, and pixels without weight.  The film stays in HBM;
 * only the byte mask (width * height) comes back, is kept by the context and is used by the passes that say adaptive = 2.  *n_resample = flagged pixels
 * (width * height when threshold <= 0: every pixel is resampled then, imagefilm.cc:917-920).  yafgpu_group_next_pass decides on the summed film of the group. */
int yafgpu_next_pass(yafgpu_ctx *ctx, const yafgpu_aa_noise_params *params, int64_t *n_resample);
int yafgpu_sync(yafgpu_ctx *ctx);
/* Film accumulators, width*height*5 floats: Pixel{Rgba col_; float weight_} (util_image_buffers.h:36-47) row-major. */
int yafgpu_read_film(yafgpu_ctx *ctx, float *rgbaw);
/* Device address of the same accumulator array (for NCCL reduction by the caller) and its float count. */
int yafgpu_film_device_ptr(yafgpu_ctx *ctx, void **d_ptr, size_t *n_floats);
/* Replace the accumulators (after a cross-GPU reduction, or to resume from a loaded film: imagefilm.cc:1506-1535). */
int yafgpu_write_film(yafgpu_ctx *ctx, const float *rgbaw);
/* Update per-pass AA multipliers between passes (integrator_tiled.cc:216-218). */
int yafgpu_set_aa_multipliers(yafgpu_ctx *ctx, float light_mult, float indirect_mult);

/* Tuning knobs that have no counterpart in the reference: "wave_size" (camera samples per wavefront wave; default: up to 2^24, limited to a quarter of the free device memory),
 * "queue_fan" (task/shadow queue capacity as a multiple of wave_size, default 8), "leaf_size" (triangles per leaf link of the BVH built by
 * the NEXT upload, 1..8), "shade_streams" (1..3 CUDA streams the per-material-type shading kernels of one iteration are spread over; default 3),
 * "debug_stack_limit" (traversal stack entries in use, only to exercise the overflow error path),
 * "ray_regeneration" (-1, default: the library decides by scene size; 0: one ray per lane and launch; 1: both traversal kernels are warp-persistent and refill idle lanes
 * from the queue; 2: only the closest-hit kernel; 3: only the shadow kernel), "regen_threshold" (1..32: refill a warp as soon as fewer lanes than this are busy; default 20),
 * "adaptive_iterations" (1, default: a wave launches as many wavefront iterations as the previous wave of the scene needed, + 1, and more only when the device
 * reports tasks still queued; 0: always the schedule's maximum, raydepth + bounces + 2),
 * "keep_nonfinite_samples" (0, default: a camera sample whose radiance is NaN/Inf is left out of the film and counted in yafgpu_stats.nonfinite_samples;
 * 1: it is splatted as the reference's ImageFilm::addSample does, imagefilm.cc, and the pixels its filter footprint covers stay NaN). */
int yafgpu_set_option(yafgpu_ctx *ctx, const char *name, long long value);
int yafgpu_get_stats(yafgpu_ctx *ctx, yafgpu_stats *out);
int yafgpu_reset_stats(yafgpu_ctx *ctx);
/* Raw CUDA stream handle of the context (cudaStream_t), so callers can time with events on the launching stream. */
void *yafgpu_stream(yafgpu_ctx *ctx);
/* Page-lock (cudaHostRegister, portable to every device) / release a host buffer the CALLER owns, so that the copies of yafgpu_upload_scene, yafgpu_trace,
 * yafgpu_read_film ... from/to it run as DMA at PCIe speed instead of through the driver's staging buffer (10^7-triangle scene: 39 ms -> 9 ms).  Optional:
 * every entry point works with pageable memory.  Registering costs about 0.1 ms per MB, so it pays for buffers that are uploaded more than once
 * (an animation re-sending the same arrays) or that a front end allocates once.  No counterpart in the reference (host-side convenience so that a caller
 * needs no CUDA runtime of its own). */
int yafgpu_host_register(void *ptr, size_t bytes);
int yafgpu_host_unregister(void *ptr);

/* ---- multi-GPU: one scene on N GPUs of one box ------------------------------------------------------------------------
 * Replaces, for a single 8-GPU node, what the reference does with worker threads pulling tiles from ImageSplitter
 * (integrator_tiled.cc:223-259, imagesplitter.cc:37-50) and, across machines, with the sum of accumulator films
 * (ImageFilm::imageFilmLoadAllInFolder, imagefilm.cc:1506-1535): the scene is replicated, member k of N renders the
 * interleaved pixel blocks of every pass that yafgpu_pass_desc.part_index / part_count describe, and after each pass the members' films are summed on the
 * first device (peer copies over NVLink + an add kernel, asynchronous to the host).  Every call mirrors the single-context call of the
 * same name; `yafgpu_group_read_film` returns the SUMMED film.  A group of one device behaves exactly like a plain context. */
typedef struct yafgpu_group yafgpu_group;
/* devices == NULL or n_devices <= 0: all GPUs of the box */
int yafgpu_group_create(yafgpu_group **out, const int *devices, int n_devices, unsigned flags);
void yafgpu_group_destroy(yafgpu_group *g);
const char *yafgpu_group_last_error(const yafgpu_group *g); /* g may be NULL: error of the last failed yafgpu_group_create */
int yafgpu_group_size(const yafgpu_group *g);
yafgpu_ctx *yafgpu_group_ctx(yafgpu_group *g, int member);  /* member context (owned by the group): per-device stats, device pointers */
int yafgpu_group_upload_scene(yafgpu_group *g, const yafgpu_scene_desc *desc);
int yafgpu_group_film_clear(yafgpu_group *g);
int yafgpu_group_write_film(yafgpu_group *g, const float *rgbaw); /* resume: the first member takes the film, the others start from zero */
int yafgpu_group_render_pass(yafgpu_group *g, const yafgpu_pass_desc *pass); /* part_index / part_count are set per member */
int yafgpu_group_read_film(yafgpu_group *g, float *rgbaw);
int yafgpu_group_next_pass(yafgpu_group *g, const yafgpu_aa_noise_params *params, int64_t *n_resample);
int yafgpu_group_sync(yafgpu_group *g); /* yafgpu_sync on every member (the first member's stream also carries the film sum) */
int yafgpu_group_set_aa_multipliers(yafgpu_group *g, float light_mult, float indirect_mult);
int yafgpu_group_set_option(yafgpu_group *g, const char *name, long long value);
int yafgpu_group_get_stats(yafgpu_group *g, yafgpu_stats *out); /* counts summed over the members, times of the slowest */

#ifdef __cplusplus
}
#endif
#endif /* YAFGPU_H */
