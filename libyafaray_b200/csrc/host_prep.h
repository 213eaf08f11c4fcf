// Host-side restatement of the reference CONSTRUCTORS: turns the parameter-level descriptors of include/yafgpu.h into
// the device records of scene_dev.cuh, computing derived constants the way the reference does (file:line per function).
// Compiled with -ffp-contract=off; uses the same YG_HD math as the kernels.
#pragma once
#include "../../include/yafgpu.h"
#include "lights.cuh"
#include <string>
#include <vector>
#include <cstring>
#include <functional>
#include <algorithm>
#include <sstream>
#include <limits>

namespace yg {



inline void cp3(float *d, const float *s) { d[0] = s[0]; d[1] = s[1]; d[2] = s[2]; }
inline void st3(float *d, V3 v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; }
inline V3 ld3(const float *s) { return mk(s[0], s[1], s[2]); }

// ShinyDiffuseMaterial ctor + factory + config(): material_shiny_diffuse.cc:26-92,599-700
inline bool prepShinyDiffuse(const yafgpu_material &m, MatDev &d, std::string &err)
{
	const yafgpu_mat_shinydiffuse &p = m.u.shinydiffuse;
	cp3(d.diffuse_col, p.color); cp3(d.mirror_col, p.mirror_color);
	d.mirror_strength = p.specular_reflect; d.transparency_strength = p.transparency; d.translucency_strength = p.translucency;
	d.diffuse_strength = p.diffuse_reflect; d.transmit_filter_strength = p.transmit_filter;
	for(int k = 0; k < 3; ++k) d.emit_col[k] = p.emit * p.color[k];
	d.emit_strength = p.emit;
	d.bsdf_flags = BsdfNone;
	if(p.emit > 0.f) d.bsdf_flags |= BsdfEmit;
	d.ior = 1.f; d.ior_squared = 1.f; d.has_fresnel = 0;
	if(p.fresnel_effect) { d.ior = p.ior; d.ior_squared = p.ior * p.ior; d.has_fresnel = 1; }
	d.use_oren_nayar = 0;
	if(p.oren_nayar)
	{	// initOrenNayar :198-204 (double)
		double s2 = (double)p.sigma * (double)p.sigma;
		d.oren_a = (float)(1.0 - 0.5 * (s2 / (s2 + 0.33)));
		d.oren_b = (float)(0.45 * s2 / (s2 + 0.09));
		d.use_oren_nayar = 1;
	}
	// config() :46-92; a bound "<component>_shader" switches the component on whatever its constant strength is
	const bool sh_mirror = d.slot[YAFGPU_SLOT_MIRROR] >= 0, sh_transp = d.slot[YAFGPU_SLOT_TRANSPARENCY] >= 0, sh_transl = d.slot[YAFGPU_SLOT_TRANSLUCENCY] >= 0;
	d.diffuse_refl_mul = 1.f;
	d.n_bsdf = 0; d.is_mirror = d.is_transparent = d.is_translucent = d.is_diffuse = 0;
	float acc = 1.f;
	if(d.mirror_strength > 0.00001f || sh_mirror)
	{
		d.is_mirror = 1;
		if(sh_mirror) {} else if(!d.has_fresnel) acc = 1.f - d.mirror_strength;
		d.bsdf_flags |= BsdfSpecular | BsdfReflect;
		d.c_flags[d.n_bsdf] = BsdfSpecular | BsdfReflect; d.c_index[d.n_bsdf] = 0; ++d.n_bsdf;
	}
	if(d.transparency_strength * acc > 0.00001f || sh_transp)
	{
		d.is_transparent = 1;
		if(!sh_transp) acc *= 1.f - d.transparency_strength;
		d.bsdf_flags |= BsdfTransmit | BsdfFilter;
		d.c_flags[d.n_bsdf] = BsdfTransmit | BsdfFilter; d.c_index[d.n_bsdf] = 1; ++d.n_bsdf;
	}
	if(d.translucency_strength * acc > 0.00001f || sh_transl)
	{
		d.is_translucent = 1;
		if(!sh_transl) acc *= 1.f - d.transparency_strength; // sic: the reference multiplies by (1 - transparency) here as well (:78)
		d.bsdf_flags |= BsdfDiffuse | BsdfTransmit;
		d.c_flags[d.n_bsdf] = BsdfDiffuse | BsdfTransmit; d.c_index[d.n_bsdf] = 2; ++d.n_bsdf;
	}
	if(d.diffuse_strength * acc > 0.00001f)
	{
		d.is_diffuse = 1;
		d.bsdf_flags |= BsdfDiffuse | BsdfReflect;
		d.c_flags[d.n_bsdf] = BsdfDiffuse | BsdfReflect; d.c_index[d.n_bsdf] = 3; ++d.n_bsdf;
	}
	// pure Lambert/Oren-Nayar surface: gets its own shading kernel (materials.cuh YG_MAT_SD_LAMBERT); same arithmetic, fewer branches
	if(d.n_bsdf == 1 && d.is_diffuse && !d.has_fresnel && d.node_count == 0) d.type = YG_MAT_SD_LAMBERT;
	(void)err;
	return true;
}

// BeerVolumeHandler ctor volumehandler_beer.cc:27-34 (the handler the glass factories create for "absorption" when the material's ParamMap
// also carries a "name": material_glass.cc:376-399, material_rough_glass.cc:382-409); log() on floats is the C double function
inline void prepBeer(const yafgpu_mat_glass &p, MatDev &d)
{
	d.absorb = 0;
	if(!p.has_absorption) return;
	const float maxlog = (float)log(1e38);
	for(int k = 0; k < 3; ++k) d.beer_sigma_a[k] = ((double)p.absorption[k] > 1e-38) ? (float)(-log((double)p.absorption[k])) : maxlog;
	const double dist = (double)p.absorption_dist;
	if(dist != 0.0) { const float f = (float)(1.0 / dist); for(int k = 0; k < 3; ++k) d.beer_sigma_a[k] *= f; }
	d.absorb = 1;
	d.bsdf_flags |= BsdfVolumetric;
}

// cauchyCoefficients__ (spectrum.cc:330-342) as the glass constructors call it (material_glass.cc:41-46, material_rough_glass.cc:40-45):
// float members, the Fraunhofer-line squares are double products narrowed to float, (ior - 1.0) is a double expression
inline void prepDispersion(const yafgpu_mat_glass &p, MatDev &d)
{
	d.disperse = 0; d.cauchy_a = 0.f; d.cauchy_b = 0.f;
	if(!(p.dispersion_power > 0.f)) return;
	d.disperse = 1;
	d.bsdf_flags |= BsdfDispersive;
	const float ior = p.ior, disp_pw = p.dispersion_power;
	const float l_f_2 = (float)(486.13 * 486.13), ld_2 = (float)(587.56 * 587.56), l_c_2 = (float)(656.27 * 656.27);
	const float vd = (float)(((double)ior - 1.0) / (double)disp_pw);   // Abbe number
	float cauchy_b = (l_c_2 - l_f_2) * vd;
	if(cauchy_b != 0.0) cauchy_b = (float)(((double)(l_f_2 * l_c_2) * ((double)ior - 1.0)) / (double)cauchy_b);
	d.cauchy_b = cauchy_b;
	d.cauchy_a = ior - cauchy_b / ld_2;
}

// GlassMaterial ctor + factory: material_glass.cc:32-49,311-379
inline bool prepGlass(const yafgpu_material &m, MatDev &d, std::string &err)
{
	const yafgpu_mat_glass &p = m.u.glass;
	d.ior = p.ior;
	// filt * filt_col + Rgb(1.f - filt): filt is a double in the factory; Rgb ops are float (double -> float at the call)
	const float filt = p.transmit_filter;
	for(int k = 0; k < 3; ++k) d.filter_col[k] = filt * p.filter_color[k] + (1.f - filt);
	cp3(d.spec_refl_col, p.mirror_color);
	d.fake_shadow = p.fake_shadows ? 1 : 0;
	d.bsdf_flags = BsdfAllSpecular;
	if(d.fake_shadow) d.bsdf_flags |= BsdfFilter;
	d.tm_flags = d.fake_shadow ? (BsdfFilter | BsdfTransmit) : (BsdfSpecular | BsdfTransmit);
	prepDispersion(p, d);
	prepBeer(p, d);
	if(d.disperse && d.absorb) { err = "glass: dispersion together with absorption is not implemented on the device (the reference applies the LAST wavelength sample's Beer transmittance to all of them, integrator_montecarlo.cc:841-845)"; return false; }
	return true;
}


inline void orenInit(double sigma, float &a, float &b)
{	// initOrenNayar (material_glossy.cc:66-72, material_coated_glossy.cc:84-90): doubles narrowed into float members
	double s2 = sigma * sigma;
	a = (float)(1.0 - 0.5 * (s2 / (s2 + 0.33)));
	b = (float)(0.45 * s2 / (s2 + 0.09));
}

// GlossyMaterial ctor + factory material_glossy.cc:32-49,407-480; CoatedGlossyMaterial material_coated_glossy.cc:41-65,465-545
inline bool prepGlossy(const yafgpu_material &m, MatDev &d, std::string &err, bool coated)
{
	const yafgpu_mat_glossy &p = m.u.glossy;
	cp3(d.gloss_col, p.color); cp3(d.diff_col, p.diffuse_color); cp3(d.gl_mirror_col, p.mirror_color);
	d.gl_exponent = p.exponent; d.gl_glossy = p.glossy_reflect; d.gl_diffuse = p.diffuse_reflect; d.as_diffuse = p.as_diffuse ? 1 : 0;
	d.gl_mirror_strength = p.specular_reflect;
	d.anisotropic = p.anisotropic ? 1 : 0; d.exp_u = p.exp_u; d.exp_v = p.exp_v;
	d.with_diffuse = p.diffuse_reflect > 0 ? 1 : 0;
	d.gl_oren_nayar = 0;
	if(p.oren_nayar) { orenInit((double)p.sigma, d.gl_oren_a, d.gl_oren_b); d.gl_oren_nayar = 1; }
	// initBsdf: p_diffuse = min(0.6, 1 - glossy / (glossy + (1 - glossy) * diffuse))
	d.p_diffuse = fminf(0.6f, 1.f - (d.gl_glossy / (d.gl_glossy + (1.f - d.gl_glossy) * d.gl_diffuse)));
	if(!coated)
	{
		d.bsdf_flags = BsdfNone;
		if(d.with_diffuse) d.bsdf_flags = BsdfDiffuse | BsdfReflect;
		d.bsdf_flags |= d.as_diffuse ? (BsdfDiffuse | BsdfReflect) : (BsdfGlossy | BsdfReflect);
	}
	else
	{
		double ior = (double)p.ior;
		if(ior == 1.0) ior = (double)1.0000001f;
		d.ior = (float)ior;
		d.c_flags[0] = BsdfSpecular | BsdfReflect;
		d.c_flags[1] = d.as_diffuse ? (BsdfDiffuse | BsdfReflect) : (BsdfGlossy | BsdfReflect);
		if(d.with_diffuse) { d.c_flags[2] = BsdfDiffuse | BsdfReflect; d.n_bsdf = 3; }
		else { d.c_flags[2] = BsdfNone; d.n_bsdf = 2; }
		d.bsdf_flags = d.c_flags[0] | d.c_flags[1] | d.c_flags[2];
	}
	(void)err;
	return true;
}

// RoughGlassMaterial ctor + factory material_rough_glass.cc:33-48,322-401
inline bool prepRoughGlass(const yafgpu_material &m, MatDev &d, std::string &err)
{
	const yafgpu_mat_glass &p = m.u.glass;
	d.ior = p.ior;
	const float filt = p.transmit_filter;
	for(int k = 0; k < 3; ++k) d.filter_col[k] = filt * p.filter_color[k] + (1.f - filt);
	cp3(d.spec_refl_col, p.mirror_color);
	d.alpha_rough = fmaxf(1e-4f, fminf(p.alpha * 0.5f, 1.f));
	d.fake_shadow = p.fake_shadows ? 1 : 0;
	d.bsdf_flags = BsdfAllGlossy;
	if(d.fake_shadow) d.bsdf_flags |= BsdfFilter;
	prepDispersion(p, d);
	prepBeer(p, d);
	if(d.disperse && d.absorb) { err = "rough_glass: dispersion together with absorption is not implemented on the device"; return false; }
	return true;
}

inline unsigned triFlagsOf(const MatDev &d)
{
	unsigned f = 0;
	if(d.visibility == YAFGPU_VIS_NORMAL || d.visibility == YAFGPU_VIS_NO_SHADOWS) f |= TRI_VIS_CAMERA;
	if(d.visibility == YAFGPU_VIS_NORMAL || d.visibility == YAFGPU_VIS_SHADOW_ONLY) f |= TRI_VIS_SHADOW;
	bool transp = false;
	if(d.type == YAFGPU_MAT_SHINYDIFFUSE || d.type == YG_MAT_SD_LAMBERT) transp = d.is_transparent; // material_shiny_diffuse.h:53
	if(d.type == YAFGPU_MAT_GLASS || d.type == YAFGPU_MAT_ROUGH_GLASS) transp = d.fake_shadow; // material_glass.h:40
	if(d.type == YAFGPU_MAT_BLEND || d.type == YAFGPU_MAT_MASK) transp = d.is_transparent; // set by finishBlend (material_blend.cc:332-335, material_mask.cc:88-91)
	if(transp) f |= TRI_TRANSPARENT;
	return f;
}

// ---- shader nodes ------------------------------------------------------------------------------------------------------
// ImageTexture after its factory, texture_image.cc:560-720 and Texture::setAdjustments texture.h:147-201
inline bool prepTexture(const yafgpu_texture &t, TexDev &d, std::string &err)
{
	memset(&d, 0, sizeof(d));
	d.type = t.type;
	if(t.ramp_items != 0)
	{	// ColorRamp::getColorInterpolated color_ramp.cc:84-164 takes std::prev of the lower bound and dereferences the lower bound itself: defined only while an item sits
		// at or in front of every position in [0, 1] and one at or behind it
		if(t.type == YAFGPU_TEX_IMAGE || t.type == YAFGPU_TEX_RGB_CUBE) { err = "a colour ramp on an image / rgb_cube texture (the reference reads none there)"; return false; }
		if(t.ramp_items < 2 || t.ramp_items > 64 || !t.ramp) { err = "colour ramp: 2..64 items"; return false; }
		for(int k = 1; k < t.ramp_items; ++k) if(!(t.ramp[5 * k + 4] >= t.ramp[5 * (k - 1) + 4])) { err = "colour ramp: items must be sorted by position"; return false; }
		if(!(t.ramp[4] <= 0.f) || !(t.ramp[5 * (t.ramp_items - 1) + 4] >= 1.f)) { err = "colour ramp: the first item must sit at position <= 0 and the last at >= 1 (the reference reads outside the ramp otherwise)"; return false; }
		if(t.ramp_mode < YAFGPU_RAMP_RGB || t.ramp_mode > YAFGPU_RAMP_HSL || t.ramp_hue_interpolation < YAFGPU_RAMPHUE_NEAR || t.ramp_hue_interpolation > YAFGPU_RAMPHUE_CCW) { err = "colour ramp: mode / hue interpolation"; return false; }
		d.n_ramp = t.ramp_items; d.ramp_mode = t.ramp_mode; d.ramp_interp = t.ramp_interpolation ? 1 : 0; d.ramp_hue = t.ramp_hue_interpolation;
	}
	if(t.type == YAFGPU_TEX_BLEND)
	{	// BlendTexture::factory texture_basic.cc:759-776: setAdjustments is called with the neutral values, whatever the parameters say
		if(t.blend_type < YAFGPU_TEXBLEND_LINEAR || t.blend_type > YAFGPU_TEXBLEND_RADIAL) { err = "blend texture: progression type"; return false; }
		d.blend_type = t.blend_type; d.flip = t.use_flip_axis ? 1 : 0;
		d.adj_intensity = d.adj_contrast = d.adj_saturation = 1.f; d.adj_mult[0] = d.adj_mult[1] = d.adj_mult[2] = 1.f;
		return true;
	}
	auto noiseOk = [](int n) { return n >= YAFGPU_NOISE_NEWPERLIN && n <= YAFGPU_NOISE_CELL; };
	const bool noise_tex = t.type == YAFGPU_TEX_CLOUDS || t.type == YAFGPU_TEX_MARBLE || t.type == YAFGPU_TEX_WOOD;
	if(t.type == YAFGPU_TEX_VORONOI)
	{	// VoronoiTexture::VoronoiTexture texture_basic.cc:391-420
		if(t.distance_metric < YAFGPU_DIST_REAL || t.distance_metric > YAFGPU_DIST_MINKOVSKY || t.color_type < 0 || t.color_type > 3) { err = "voronoi texture: distance_metric / colour type"; return false; }
		d.size = t.noise_size; d.mk_exp = t.mk_exponent; d.dist_metric = t.distance_metric; d.color_type = t.color_type;
		for(int k = 0; k < 4; ++k) { d.vw[k] = t.vor_weight[k]; d.vaw[k] = std::fabs(t.vor_weight[k]); }
		d.iscale = t.intensity_scale;
	}
	else if(t.type == YAFGPU_TEX_MUSGRAVE)
	{	// MusgraveTexture::MusgraveTexture texture_basic.cc:509-527
		if(!noiseOk(t.noise_type) || t.musgrave_type < YAFGPU_MUSGRAVE_FBM || t.musgrave_type > YAFGPU_MUSGRAVE_RIDGEDMF) { err = "musgrave texture: noise_type / musgrave_type"; return false; }
		if(!(t.mus_octaves >= 0.f && t.mus_octaves <= 64.f)) { err = "musgrave texture: octaves must be 0..64"; return false; }
		d.noise_type = t.noise_type; d.mus_type = t.musgrave_type; d.size = t.noise_size; d.iscale = t.intensity_scale;
		d.mus_h = t.mus_h; d.mus_lac = t.mus_lacunarity; d.mus_oct = t.mus_octaves; d.mus_offset = t.mus_offset; d.mus_gain = t.mus_gain;
		for(int k = 0; k < 3; ++k) { d.color1[k] = t.color1[k]; d.color2[k] = t.color2[k]; }
	}
	else if(t.type == YAFGPU_TEX_DISTORTED_NOISE)
	{	// DistortedNoiseTexture::DistortedNoiseTexture texture_basic.cc:602-610
		if(!noiseOk(t.noise_type) || !noiseOk(t.noise_type2)) { err = "distorted_noise texture: noise_type1 / noise_type2"; return false; }
		d.noise_type = t.noise_type; d.noise_type2 = t.noise_type2; d.distort = t.distort; d.size = t.noise_size;
		for(int k = 0; k < 3; ++k) { d.color1[k] = t.color1[k]; d.color2[k] = t.color2[k]; }
	}
	else if(noise_tex)
	{	// CloudsTexture / MarbleTexture / WoodTexture after their constructors (texture_basic.cc:96-107,177-188,257-266)
		if(t.noise_type < YAFGPU_NOISE_NEWPERLIN || t.noise_type > YAFGPU_NOISE_CELL) { err = "noise texture: noise_type"; return false; }
		if(t.noise_depth < 0 || t.noise_depth > 29) { err = "noise texture: depth must be 0..29 (turbulence__ shifts 1 << (depth + 1))"; return false; }
		d.noise_type = t.noise_type; d.depth = t.noise_depth; d.size = t.noise_size; d.hard = t.noise_hard ? 1 : 0;
		for(int k = 0; k < 3; ++k) { d.color1[k] = t.color1[k]; d.color2[k] = t.color2[k]; }
		d.bias = t.type == YAFGPU_TEX_CLOUDS ? t.clouds_bias : 0; d.turb = t.turbulence; d.sharpness = t.sharpness;
		d.shape = t.wave_shape; d.rings = t.wood_rings ? 1 : 0;
		if(d.bias < 0 || d.bias > 2 || d.shape < YAFGPU_WAVE_SIN || d.shape > YAFGPU_WAVE_TRI) { err = "noise texture: bias / shape"; return false; }
	}
	else if(t.type != YAFGPU_TEX_IMAGE && t.type != YAFGPU_TEX_RGB_CUBE) { err = "texture type " + std::to_string(t.type) + " is not implemented on the device"; return false; }
	if(t.type == YAFGPU_TEX_IMAGE)
	if(!t.rgba || t.width <= 0 || t.height <= 0) { err = "image texture without texels"; return false; }
	if(t.type == YAFGPU_TEX_IMAGE && (t.interpolate < YAFGPU_TEXINTERP_NONE || t.interpolate > YAFGPU_TEXINTERP_BICUBIC)) { err = "texture interpolation (mipmap_trilinear / mipmap_ewa need ray differentials) is not implemented on the device"; return false; }
	if(t.type == YAFGPU_TEX_IMAGE && (t.clip_mode < YAFGPU_TEXCLIP_EXTEND || t.clip_mode > YAFGPU_TEXCLIP_CHECKER)) { err = "texture clip mode"; return false; }
	d.w = t.width; d.h = t.height; d.interp = t.interpolate; d.clip = t.clip_mode;
	d.xrepeat = t.xrepeat; d.yrepeat = t.yrepeat; d.mirror_x = t.mirror_x ? 1 : 0; d.mirror_y = t.mirror_y ? 1 : 0; d.rot90 = t.rot90 ? 1 : 0;
	// setCrop :224-229
	d.cropminx = t.cropmin_x; d.cropmaxx = t.cropmax_x; d.cropminy = t.cropmin_y; d.cropmaxy = t.cropmax_y;
	d.cropx = ((double)d.cropminx != 0.0) || ((double)d.cropmaxx != 1.0);
	d.cropy = ((double)d.cropminy != 0.0) || ((double)d.cropmaxy != 1.0);
	d.checker_even = t.checker_even ? 1 : 0; d.checker_odd = t.checker_odd ? 1 : 0; d.checker_dist = t.checker_dist;
	d.normalmap = t.normalmap ? 1 : 0; d.color_space = t.color_space; d.gamma = t.gamma;
	d.adj_intensity = t.adj_intensity; d.adj_contrast = t.adj_contrast; d.adj_saturation = t.adj_saturation; d.adj_hue = t.adj_hue;
	d.adj_clamp = t.adj_clamp ? 1 : 0; d.adj_mult[0] = t.adj_mult_red; d.adj_mult[1] = t.adj_mult_green; d.adj_mult[2] = t.adj_mult_blue;
	d.adj_set = (t.adj_intensity != 1.f || t.adj_contrast != 1.f || t.adj_saturation != 1.f || t.adj_hue != 0.f || t.adj_mult_red != 1.f ||
	             t.adj_mult_green != 1.f || t.adj_mult_blue != 1.f || t.adj_clamp) ? 1 : 0;
	return true;
}

// NodeMaterial::loadNodes / solveNodesOrder / getNodeList (material_node.cc:46-119,140-214) for one material: appends the material's
// reachable nodes to `out` in a dependency order (the reference's order is one particular depth-first order; node results do not depend
// on which topological order is used), and fills the node fields of `d`.
inline bool prepNodes(const yafgpu_material &m, const yafgpu_scene_desc &sd, const std::vector<TexDev> &tex, MatDev &d, std::vector<NodeDev> &out, std::string &err)
{
	for(int s = 0; s < YAFGPU_N_SLOTS; ++s) d.slot[s] = -1;
	d.node_first = (int)out.size(); d.node_count = 0; d.color_mask = d.bump_mask = 0u;
	const int n = m.node_count;
	if(n <= 0) return true;
	if(m.node_first < 0 || (size_t)m.node_first + (size_t)n > sd.n_nodes || !sd.nodes) { err = "shader node range out of bounds"; return false; }
	const yafgpu_shader_node *nd = sd.nodes + m.node_first;
	auto deps = [&](int i, int *dp) { int k = 0; const yafgpu_shader_node &q = nd[i];
		if(q.type == YAFGPU_NODE_MIX) { if(q.input1 >= 0) dp[k++] = q.input1; if(q.input2 >= 0) dp[k++] = q.input2; if(q.factor >= 0) dp[k++] = q.factor; }
		else if(q.type == YAFGPU_NODE_LAYER) { if(q.input1 >= 0) dp[k++] = q.input1; if(q.input2 >= 0) dp[k++] = q.input2; }
		return k; };
	for(int i = 0; i < n; ++i)
	{
		int dp[3]; const int k = deps(i, dp);
		for(int j = 0; j < k; ++j) if(dp[j] >= n) { err = "shader node input index out of range"; return false; }
		if(nd[i].type == YAFGPU_NODE_LAYER && nd[i].input1 < 0) { err = "layer node without input"; return false; }
		if(nd[i].type < YAFGPU_NODE_TEXTURE_MAPPER || nd[i].type > YAFGPU_NODE_LAYER) { err = "shader node type " + std::to_string(nd[i].type) + " is not implemented on the device"; return false; }
	}
	std::vector<int> pos(n, -1), state(n, 0), order;
	bool cyc = false;
	std::function<void(int)> visit = [&](int i) { // recursiveSolver__ :46-58
		if(state[i] == 2) return;
		if(state[i] == 1) { cyc = true; return; }
		state[i] = 1;
		int dp[3]; const int k = deps(i, dp);
		for(int j = 0; j < k; ++j) visit(dp[j]);
		state[i] = 2; pos[i] = (int)order.size(); order.push_back(i); };
	for(int s = 0; s < YAFGPU_N_SLOTS; ++s)
	{
		if(m.shader[s] < 0) continue;
		if(m.shader[s] >= n) { err = "shader slot index out of range"; return false; }
		visit(m.shader[s]);
	}
	if(cyc) { err = "shader node graph has a cycle"; return false; }
	if(order.empty()) return true; // nodes declared but none bound: nothing is ever evaluated
	if(order.size() > YG_MAX_NODES) { err = "more than " + std::to_string(YG_MAX_NODES) + " shader nodes in one material"; return false; }
	std::function<void(int, unsigned &)> mark = [&](int i, unsigned &mask) { // recursiveFinder__ :60-72
		if(mask >> pos[i] & 1u) return;
		mask |= 1u << pos[i];
		int dp[3]; const int k = deps(i, dp);
		for(int j = 0; j < k; ++j) mark(dp[j], mask); };
	for(int s = 0; s < YAFGPU_N_SLOTS; ++s)
	{
		if(m.shader[s] < 0) continue;
		d.slot[s] = pos[m.shader[s]];
		mark(m.shader[s], s == YAFGPU_SLOT_BUMP ? d.bump_mask : d.color_mask);
	}
	d.node_count = (int)order.size();
	for(int i : order)
	{
		const yafgpu_shader_node &q = nd[i];
		NodeDev o; memset(&o, 0, sizeof(o));
		o.type = q.type; o.mode = q.mode;
		o.in1 = q.input1 >= 0 ? pos[q.input1] : -1; o.in2 = q.input2 >= 0 ? pos[q.input2] : -1; o.fac = q.factor >= 0 ? pos[q.factor] : -1;
		for(int k = 0; k < 4; ++k) { o.col[k] = q.color[k]; o.col2[k] = q.color2[k]; }
		o.f1 = q.scalar; o.f2 = q.scalar2;
		if(q.type == YAFGPU_NODE_TEXTURE_MAPPER)
		{	// TextureMapperNode::factory :344-420 and setup() :33-55
			if(q.texture < 0 || (size_t)q.texture >= tex.size()) { err = "texture_mapper: texture index out of range"; return false; }
			const TexDev &t = tex[q.texture];
			o.tex = q.texture; o.coords = q.texco; o.discrete = t.type == YAFGPU_TEX_IMAGE;
			o.proj = o.discrete ? q.mapping : YAFGPU_TEXPROJ_PLAIN; // the factory honours "mapping" for discrete() textures only (shader_node_basic.cc:381)
			if(q.texco < 0 || q.texco > YAFGPU_TEXCO_TANGENT || q.mapping < 0 || q.mapping > YAFGPU_TEXPROJ_SPHERE) { err = "texture_mapper: texco / mapping"; return false; }
			o.map_x = std::min(3, std::max(0, q.proj_x)); o.map_y = std::min(3, std::max(0, q.proj_y)); o.map_z = std::min(3, std::max(0, q.proj_z));
			for(int k = 0; k < 3; ++k) { o.scale[k] = q.scale[k]; o.offset[k] = 2 * q.offset[k]; }
			o.do_scalar = q.do_scalar ? 1 : 0;
			for(int k = 0; k < 12; ++k) o.mtx[k] = q.transform[k];
			if(o.discrete) { o.d_u = 1.f / (float)t.w; o.d_v = 1.f / (float)t.h; }
			else o.d_u = o.d_v = 0.0002f;
			o.bump_str = q.bump_strength / length(mk(o.scale[0], o.scale[1], o.scale[2]));
			if(!t.normalmap) o.bump_str /= 100.0f;
		}
		else if(q.type == YAFGPU_NODE_LAYER)
		{	// LayerNode::factory :203-244 (colfac / valfac / def_val are read as doubles and narrowed by the constructor)
			o.colfac = q.colfac; o.valfac = q.valfac;
			o.texflag = (q.no_rgb ? 1u : 0u) | (q.stencil ? 2u : 0u) | (q.negative ? 4u : 0u) | (q.use_alpha ? 8u : 0u);
			o.do_color = q.do_color ? 1 : 0; o.do_scalar_l = q.layer_do_scalar ? 1 : 0; o.color_input = q.color_input ? 1 : 0; o.use_alpha = q.use_alpha ? 1 : 0;
			o.col[3] = 1.f; // Rgba(def_col)
		}
		out.push_back(o);
	}
	return true;
}

inline bool prepMaterial(const yafgpu_material &m, MatDev &d, std::string &err)
{
	// node fields (node_first .. slot[]) were filled by prepNodes and are kept
	{ MatDev keep = d; memset(&d, 0, sizeof(d)); d.node_first = keep.node_first; d.node_count = keep.node_count; d.color_mask = keep.color_mask; d.bump_mask = keep.bump_mask; memcpy(d.slot, keep.slot, sizeof(d.slot)); }
	d.diffuse_refl_mul = 1.f;
	if(d.node_count > 0)
	{	// which "<slot>_shader" parameters each factory looks up (material_*.cc node_list[...]); a slot the type does not have is ignored there
		static const unsigned allowed[9] = {
			/* shinydiffuse */ 1u << YAFGPU_SLOT_DIFFUSE | 1u << YAFGPU_SLOT_MIRROR_COLOR | 1u << YAFGPU_SLOT_BUMP | 1u << YAFGPU_SLOT_MIRROR | 1u << YAFGPU_SLOT_TRANSPARENCY |
			                   1u << YAFGPU_SLOT_TRANSLUCENCY | 1u << YAFGPU_SLOT_SIGMA_OREN | 1u << YAFGPU_SLOT_DIFFUSE_REFL | 1u << YAFGPU_SLOT_IOR | 1u << YAFGPU_SLOT_WIREFRAME,
			/* glossy */       1u << YAFGPU_SLOT_DIFFUSE | 1u << YAFGPU_SLOT_GLOSSY | 1u << YAFGPU_SLOT_GLOSSY_REFLECT | 1u << YAFGPU_SLOT_BUMP | 1u << YAFGPU_SLOT_SIGMA_OREN |
			                   1u << YAFGPU_SLOT_EXPONENT | 1u << YAFGPU_SLOT_WIREFRAME | 1u << YAFGPU_SLOT_DIFFUSE_REFL,
			/* coated */       1u << YAFGPU_SLOT_DIFFUSE | 1u << YAFGPU_SLOT_GLOSSY | 1u << YAFGPU_SLOT_GLOSSY_REFLECT | 1u << YAFGPU_SLOT_BUMP | 1u << YAFGPU_SLOT_SIGMA_OREN |
			                   1u << YAFGPU_SLOT_EXPONENT | 1u << YAFGPU_SLOT_WIREFRAME | 1u << YAFGPU_SLOT_DIFFUSE_REFL | 1u << YAFGPU_SLOT_MIRROR | 1u << YAFGPU_SLOT_MIRROR_COLOR | 1u << YAFGPU_SLOT_IOR,
			/* glass */        1u << YAFGPU_SLOT_MIRROR_COLOR | 1u << YAFGPU_SLOT_BUMP | 1u << YAFGPU_SLOT_FILTER_COLOR | 1u << YAFGPU_SLOT_IOR | 1u << YAFGPU_SLOT_WIREFRAME,
			/* rough glass */  1u << YAFGPU_SLOT_MIRROR_COLOR | 1u << YAFGPU_SLOT_BUMP | 1u << YAFGPU_SLOT_FILTER_COLOR | 1u << YAFGPU_SLOT_IOR | 1u << YAFGPU_SLOT_WIREFRAME | 1u << YAFGPU_SLOT_ROUGHNESS,
			/* blend */        1u << YAFGPU_SLOT_BLEND,
			0u, 0u, 0u };
		const unsigned allow = m.type == YAFGPU_MAT_MASK ? 1u << YAFGPU_SLOT_BLEND : (m.type < 0 || m.type > 8) ? 0u : allowed[m.type];
		if(allow == 0u) { err = "shader nodes on this material type are not implemented on the device"; return false; }
		for(int s = 0; s < YAFGPU_N_SLOTS; ++s) if(d.slot[s] >= 0 && !(allow >> s & 1u)) { err = "shader slot " + std::to_string(s) + " does not exist on material type " + std::to_string(m.type); return false; }
		if(d.slot[YAFGPU_SLOT_WIREFRAME] >= 0) { err = "wireframe_shader is not implemented on the device"; return false; }
	}
	d.type = m.type; d.visibility = m.visibility; d.receive_shadows = m.receive_shadows; d.flat = m.flat_material;
	d.additional_depth = m.additional_depth; d.transp_bias_factor = m.transparent_bias_factor;
	d.transp_bias_mult = m.transparent_bias_multiply_raydepth; d.sampling_factor = m.sampling_factor;
	switch(m.type)
	{
		case YAFGPU_MAT_SHINYDIFFUSE: return prepShinyDiffuse(m, d, err);
		case YAFGPU_MAT_GLASS: return prepGlass(m, d, err);
		case YAFGPU_MAT_LIGHT: // LightMaterial material_simple.cc:38-41,63-73
			for(int k = 0; k < 3; ++k) d.simple_col[k] = m.u.simple.color[k] * m.u.simple.strength;
			d.double_sided = m.u.simple.double_sided; d.bsdf_flags = BsdfEmit;
			return true;
		case YAFGPU_MAT_NULL: d.bsdf_flags = BsdfNone; return true;
		case YAFGPU_MAT_GLOSSY: return prepGlossy(m, d, err, false);
		case YAFGPU_MAT_COATED_GLOSSY: return prepGlossy(m, d, err, true);
		case YAFGPU_MAT_ROUGH_GLASS: return prepRoughGlass(m, d, err);
		case YAFGPU_MAT_MIRROR: // MirrorMaterial material_glass.h:79-84
			for(int k = 0; k < 3; ++k) d.simple_col[k] = m.u.simple.color[k] * m.u.simple.strength;
			d.bsdf_flags = BsdfSpecular;
			return true;
		case YAFGPU_MAT_MASK: // MaskMaterial ctor + factory material_mask.cc:30-37,150-198; flags resolved by finishBlend() once all materials exist
			d.mat1 = m.u.blend.material1; d.mat2 = m.u.blend.material2; d.blend_val = (float)(double)m.u.blend.blend_value;
			if(d.slot[YAFGPU_SLOT_BLEND] < 0) { err = "mask_mat without a \"mask\" shader node"; return false; }
			return true;
		case YAFGPU_MAT_BLEND: // BlendMaterial ctor material_blend.cc:37-49; flags/additional depth resolved by finishBlend() once all materials exist
			d.mat1 = m.u.blend.material1; d.mat2 = m.u.blend.material2; d.blend_val = m.u.blend.blend_value;
			return true;
		default:
			err = "material type " + std::to_string(m.type) + " is not implemented on the device yet";
			return false;
	}
}

inline bool matIsTransparent(const MatDev &d)
{
	if(d.type == YAFGPU_MAT_SHINYDIFFUSE || d.type == YG_MAT_SD_LAMBERT) return d.is_transparent;
	if(d.type == YAFGPU_MAT_GLASS || d.type == YAFGPU_MAT_ROUGH_GLASS) return d.fake_shadow;
	return false;
}
// second pass over blend materials (children must be leaf materials)
inline bool finishBlend(std::vector<MatDev> &mats, size_t i, std::string &err)
{
	MatDev &d = mats[i];
	if(d.mat1 < 0 || d.mat2 < 0 || (size_t)d.mat1 >= mats.size() || (size_t)d.mat2 >= mats.size()) { err = "blend_mat: child index out of range"; return false; }
	const MatDev &a = mats[d.mat1], &b = mats[d.mat2];
	if(d.type == YAFGPU_MAT_MASK)
	{	// the chosen child's record replaces the mask's per hit (nodes.cuh), so children are plain materials (a blend_mat of plain materials included)
		if(a.type == YAFGPU_MAT_MASK || b.type == YAFGPU_MAT_MASK || a.node_count > 0 || b.node_count > 0) { err = "mask_mat of materials with shader nodes is not implemented on the device"; return false; }
		d.bsdf_flags = a.bsdf_flags | b.bsdf_flags;
		d.is_transparent = (matIsTransparent(a) || matIsTransparent(b)) ? 1 : 0;
		return true;
	}
	if(a.type == YAFGPU_MAT_MASK || b.type == YAFGPU_MAT_MASK) { err = "blend_mat of mask_mat is not implemented on the device"; return false; }
	if(a.type == YAFGPU_MAT_BLEND || b.type == YAFGPU_MAT_BLEND) { err = "blend_mat of blend_mat is not implemented on the device"; return false; }
	d.bsdf_flags = a.bsdf_flags | b.bsdf_flags;
	d.additional_depth = a.additional_depth > b.additional_depth ? a.additional_depth : b.additional_depth;
	d.is_transparent = (matIsTransparent(a) || matIsTransparent(b)) ? 1 : 0;
	if(a.absorb || b.absorb) { err = "blend_mat of a glass with Beer absorption is not implemented on the device"; return false; }
	return true;
}

// Light constructors
inline bool prepLight(const yafgpu_light &l, LightDev &d, std::string &err)
{
	memset(&d, 0, sizeof(d));
	d.type = l.type; d.enabled = l.light_enabled; d.cast_shadows = l.cast_shadows; d.photon_only = l.photon_only;
	d.samples = l.samples; d.object = l.object; d.double_sided = l.double_sided; d.abs_intersect = l.abs_intersect;
	d.clamp_intersect = l.clamp_intersect; d.shoot_caustic = l.shoot_caustic;
	d.dirac = l.type == YAFGPU_LIGHT_POINT; d.can_intersect = l.type != YAFGPU_LIGHT_POINT; // sun, area, mesh (tree built), background
	switch(l.type)
	{
		case YAFGPU_LIGHT_DIRECTIONAL: // DirectionalLight::DirectionalLight light_directional.cc:31-42 (+ factory :112-158)
		{
			for(int k = 0; k < 3; ++k) d.color[k] = l.color[k] * l.power;
			st3(d.fnormal, normalize(ld3(l.point1)));
			cp3(d.pos, l.from); d.area = l.radius; d.double_sided = !l.infinite;
			d.dirac = 1; d.can_intersect = 0;
			return true;
		}
		case YAFGPU_LIGHT_SPOT: // SpotLight::SpotLight light_spot.cc:28-69
		{
			const V3 from = ld3(l.from), to = ld3(l.point1);
			const V3 ndir = normalize(from - to), dir = -ndir;
			for(int k = 0; k < 3; ++k) d.color[k] = l.color[k] * l.power;
			V3 du, dv;
			createCs(dir, du, dv);
			const double rad_angle = (double)l.cone_angle * 0.01745329251994329576922;
			const double rad_inner_angle = rad_angle * (double)(1.f - l.blend);
			d.pdf = fCos((float)rad_inner_angle);   // cos_start_
			d.cos_angle = fCos((float)rad_angle);   // cos_end_
			d.invpdf = (float)(1.0 / (double)(d.pdf - d.cos_angle)); // icos_diff_
			d.area = l.shadow_fuzzy;
			cp3(d.pos, l.from); st3(d.fnormal, ndir); st3(d.normal, dir); st3(d.du, du); st3(d.dv, dv);
			d.dirac = !l.soft_shadows; d.can_intersect = l.soft_shadows != 0;
			return true;
		}
		case YAFGPU_LIGHT_IES: // finished by buildIesTables (needs the table vector)
			d.dirac = !l.soft_shadows; d.can_intersect = 0;
			return true;
		case YAFGPU_LIGHT_SPHERE: // SphereLight::SphereLight light_sphere.cc:31-41
		{
			if(l.object > 0) { err = "spherelight linked to a sphere primitive is not supported"; return false; }
			for(int k = 0; k < 3; ++k) d.color[k] = l.color[k] * l.power;
			cp3(d.pos, l.from);
			d.area = l.radius * l.radius;                           // square_radius_
			d.inv_area = (float)((double)d.area * 1.000003815);     // square_radius_epsilon_
			d.dirac = 0; d.can_intersect = 0;
			return true;
		}
		case YAFGPU_LIGHT_POINT: // PointLight::PointLight light_point.cc:29-36
			for(int k = 0; k < 3; ++k) d.color[k] = l.color[k] * l.power;
			cp3(d.pos, l.from);
			return true;
		case YAFGPU_LIGHT_AREA: // AreaLight::AreaLight light_area.cc:34-53; factory passes (p1 - corner), (p2 - corner) :200
		{
			V3 c = ld3(l.from), v1 = ld3(l.point1) - c, v2 = ld3(l.point2) - c;
			V3 fn = cross(v2, v1);
			// color_ = col * inte * M_PI : Rgb*float, then Rgb * (float)M_PI (Rgb::operator*(float))
			for(int k = 0; k < 3; ++k) d.color[k] = (l.color[k] * l.power) * (float)YG_M_PI;
			d.area = normLen(fn);
			d.inv_area = (float)(1.0 / (double)d.area);
			V3 nrm = -fn;
			V3 du = normalize(v1);
			V3 dv = cross(nrm, du);
			st3(d.pos, c); st3(d.to_x, v1); st3(d.to_y, v2); st3(d.fnormal, fn); st3(d.normal, nrm);
			st3(d.du, du); st3(d.dv, dv);
			st3(d.c2, c + v1); st3(d.c3, c + (v1 + v2)); st3(d.c4, c + v2);
			return true;
		}
		case YAFGPU_LIGHT_SUN: // SunLight::SunLight light_sun.cc:28-42
		{
			for(int k = 0; k < 3; ++k) d.color[k] = l.color[k] * l.power;
			V3 dir = ld3(l.from);
			V3 ndir = normalize(dir);
			V3 du, dv;
			createCs(dir, du, dv); // sic: the un-normalised parameter
			float angle = l.angle;
			if(angle > 80.f) angle = 80.f;
			d.cos_angle = fCos((float)((double)angle * 0.01745329251994329576922));
			d.invpdf = (float)(YG_M_2PI * (double)(1.f - d.cos_angle));
			d.pdf = (float)(1.0 / (double)d.invpdf);
			for(int k = 0; k < 3; ++k) d.col_pdf[k] = d.color[k] * d.pdf;
			st3(d.pos, ndir); st3(d.du, du); st3(d.dv, dv);
			return true;
		}
		case YAFGPU_LIGHT_MESH: // MeshLight::factory light_meshlight.cc:251: color * (float)power * M_PI; tables/BVH built by the caller (initIs)
			for(int k = 0; k < 3; ++k) d.color[k] = (l.color[k] * l.power) * (float)YG_M_PI;
			d.accel = -1;
			return true;
		case YAFGPU_LIGHT_BGPORTAL: // BackgroundPortalLight ctor + factory light_background_portal.cc:34-44,236-266; tables/BVH as for the mesh light (initIs :58-78)
			d.power = l.power; d.accel = -1; d.double_sided = 0; d.clamp_intersect = 0.f;
			return true;
		case YAFGPU_LIGHT_BACKGROUND: // BackgroundLight ctor light_background.cc:56-65; tables built by buildBgLightTables (init)
			return true;
		default:
			err = "light type " + std::to_string(l.type) + " is not supported";
			return false;
	}
}


// MeshLight::initIs light_meshlight.cc:55-73 + Pdf1D/cumulateStep1DDf__ util_sample.h:89-124: appends cdf[count+1] to `tab`
// IesLight::IesLight (light_ies.cc:30-50) + IesData::parseIesFile (util_ies.h:131-379) on the file CONTENTS.  The parser is the reference's token
// stream, restated with the same iostream extractions (so malformed numbers behave alike); its console prints are dropped.  Differences, all on
// input the reference mishandles: a file without any "TILT=" token loops forever there and is rejected here; and getRadiance's interval search reads
// one element past each angle map (util_ies.h:94-111).  That element is allocator slack or the next chunk's header (zero or a denormal), never
// greater than an angle that has reached the last map entry, so for such an angle no interval matches, the index stays 0 and the reference
// EXTRAPOLATES from the first cell (often to a negative radiance, which doLightEstimation then drops through its pdf test).  The sampled direction of
// a soft-shadow ieslight is not unit length, so this happens routinely (world z >= 1 and y < 0 give a horizontal angle of exactly 360).
// The horizontal map carries -INF in that slot, which reproduces exactly that (pinned by the soft-shadow parity renders).  The vertical map can only
// be overrun in the thin ring where the fast-cosine cut-off lets the angle pass its last entry; there the reference's result depends on heap contents
// (its golden renders behave as if the element were large: the last cell is kept), so that slot carries +INF.
// IESNA LM-63 photometric file -> the three tables IesData holds (vertical angles, horizontal angles, candela map).  Own reader: the
// text is cut into whitespace-separated tokens once and the LM-63 record is read field by field from a cursor; the numbers it yields are
// what the reference's stream extraction yields on a well-formed file (pinned by the IES parity renders), the code shares nothing with it.
struct IesTokens
{
	std::vector<std::pair<const char *, size_t>> tok; size_t pos = 0;
	explicit IesTokens(const char *p, size_t n)
	{
		size_t i = 0;
		while(i < n)
		{
			while(i < n && isspace((unsigned char)p[i])) ++i;
			const size_t b = i;
			while(i < n && !isspace((unsigned char)p[i])) ++i;
			if(i > b) tok.emplace_back(p + b, i - b);
		}
	}
	bool done() const { return pos >= tok.size(); }
	std::string text(size_t k) const { return std::string(tok[k].first, tok[k].second); }
	void skip(size_t k) { pos += k; }
	double number() { if(done()) return 0.0; const std::string t = text(pos++); return strtod(t.c_str(), nullptr); }
	bool seek(const char *needle)   // position the cursor just behind the first token that contains `needle`
	{
		for(size_t k = pos; k < tok.size(); ++k) if(text(k).find(needle) != std::string::npos) { pos = k + 1; return true; }
		return false;
	}
};

inline bool buildIesTables(const yafgpu_light &l, LightDev &d, std::vector<float> &tab, std::string &err)
{
	if(!l.ies_data || l.ies_size < 7) { err = "ieslight: IES file missing or too small"; return false; }
	if(strncmp(l.ies_data, "IESNA:", 6) != 0) { err = "ieslight: wrong file format, first characters are not \"IESNA:\""; return false; }
	IesTokens in(l.ies_data, l.ies_size);
	// keyword lines end at the TILT= line; TILT=INCLUDE carries <lamp-to-luminaire geometry> <n pairs> <n angles> <n multiplying factors>
	if(!in.seek("TILT=")) { err = "ieslight: no TILT= entry in the IES file"; return false; }
	if(in.text(in.pos - 1) == "TILT=INCLUDE")
	{
		in.skip(1);
		const int n_pairs = (int)in.number();
		in.skip((size_t)std::max(0, n_pairs) * 2);
	}
	// photometric record: lamps, lumens per lamp, candela multiplier, vertical count, horizontal count, photometric type, units type,
	// width, length, height, ballast factor, (future use), input watts
	in.skip(2);
	(void)in.number();                       // candela multiplier: read, not applied (IesData keeps the raw candela values)
	int nv = (int)in.number(), nh = (int)in.number();
	const int type = (int)in.number();
	in.skip(1 + 3 + 3);
	if(nv < 1 || nh < 1 || nv > (1 << 16) || nh > (1 << 16)) { err = "ieslight: implausible angle counts in the IES file"; return false; }
	std::vector<float> vert((size_t)nv);
	for(float &a : vert) a = (float)in.number();
	float max_v = *std::max_element(vert.begin(), vert.end());
	if(max_v < 0.f) max_v = 0.f;
	if(vert[0] > 0.f)
	{	// a vertical range that does not start at 0 is shifted down to 0
		const float first = vert[0];
		max_v -= first;
		for(float &a : vert) a -= first;
	}
	max_v = (float)((double)max_v * 0.01745329251994329576922); // degrees -> radians, float * double as in IesData
	// a type C file with one horizontal plane is axially symmetric: the plane is duplicated at 180 degrees
	const bool mirrored = type == 1 && nh == 1;
	const int nh_file = nh;
	if(mirrored) nh = 2;
	std::vector<float> hor((size_t)nh), rad((size_t)nh * nv);
	for(int i = 0; i < nh_file; ++i) hor[i] = (float)in.number();
	for(int i = 0; i < nh_file; ++i) for(int j = 0; j < nv; ++j) rad[(size_t)i * nv + j] = (float)in.number();
	if(mirrored) { hor[1] = 180.f; std::copy(rad.begin(), rad.begin() + nv, rad.begin() + nv); }
	float max_rad = 0.f;
	for(float r : rad) if(max_rad < r) max_rad = r;
	max_rad = 1.f / max_rad;
	d.cdf_offset = (int)tab.size(); d.cdf_count = nv + 1 + nh + 1 + nh * nv;
	d.ies_nv = nv; d.ies_nh = nh; d.ies_type = type; d.ies_max_rad = max_rad;
	tab.insert(tab.end(), vert.begin(), vert.end()); tab.push_back(std::numeric_limits<float>::infinity());
	tab.insert(tab.end(), hor.begin(), hor.end()); tab.push_back(-std::numeric_limits<float>::infinity());
	tab.insert(tab.end(), rad.begin(), rad.end());
	// constructor: ndir_ = (from - to).normalize(), dir_ = -ndir_, createCs__(dir_, du_, dv_), cos_end_ = fCos__(max vertical angle), color_ = col * power
	const V3 from = ld3(l.from), to = ld3(l.point1);
	const V3 ndir = normalize(from - to), dir = -ndir;
	V3 du, dv;
	createCs(dir, du, dv);
	d.cos_angle = fCos(max_v);
	for(int k = 0; k < 3; ++k) d.color[k] = l.color[k] * l.power;
	cp3(d.pos, l.from); st3(d.fnormal, ndir); st3(d.normal, dir); st3(d.du, du); st3(d.dv, dv);
	d.dirac = !l.soft_shadows; d.can_intersect = 0;
	return true;
}

inline void buildMeshLightTables(const yafgpu_geometry &g, const yafgpu_object &o, LightDev &d, std::vector<float> &tab)
{
	const int n = o.n_tris;
	std::vector<float> areas(n);
	double total_area = 0.0;
	for(int i = 0; i < n; ++i)
	{
		const int *vi = g.tri_vidx + 3 * (size_t)(o.first_tri + i);
		const V3 a = ld3(g.points + 3 * (size_t)vi[0]), b = ld3(g.points + 3 * (size_t)vi[1]), c = ld3(g.points + 3 * (size_t)vi[2]);
		areas[i] = (float)(0.5 * (double)length(cross(b - a, c - a))); // Triangle::surfaceArea triangle.cc:169-179
		total_area += (double)areas[i];
	}
	d.area = (float)total_area; d.inv_area = (float)(1.0 / total_area);
	d.tri_first = o.first_tri; d.tri_count = n;
	d.cdf_offset = (int)tab.size(); d.cdf_count = n;
	std::vector<float> cdf(n + 1);
	double c = 0.0, delta = 1.0 / (double)n;
	cdf[0] = 0.f;
	for(int i = 1; i < n + 1; ++i) { c += (double)areas[i - 1] * delta; cdf[i] = (float)c; }
	const float integral = (float)c;
	for(int i = 1; i < n + 1; ++i) cdf[i] /= integral;
	d.cdf_integral = integral;
	tab.insert(tab.end(), cdf.begin(), cdf.end());
}

// BackgroundLight::init light_background.cc:78-120 (layout read by lights.cuh bgRow/bgCalcFrom*)
inline void buildBgLightTables(const BackgroundDev &bg, LightDev &d, std::vector<float> &tab)
{
	const int nv = 360, max_u = 720, min_s = 16;
	std::vector<float> fu(max_u), fv(nv);
	struct Row { std::vector<float> func, cdf; float integral; };
	std::vector<Row> rows(nv);
	auto cumulate = [](const std::vector<float> &f, int n, std::vector<float> &cdf, float &integral)
	{
		cdf.assign(n + 1, 0.f);
		double c = 0.0, delta = 1.0 / (double)n;
		for(int i = 1; i < n + 1; ++i) { c += (double)f[i - 1] * delta; cdf[i] = (float)c; }
		integral = (float)c;
		for(int i = 1; i < n + 1; ++i) cdf[i] /= integral;
	};
	const float inv = 1.f / (float)nv;
	for(int y = 0; y < nv; ++y)
	{
		const float fy = ((float)y + 0.5f) * inv;
		const float sintheta = bgSinSample(fy);
		const int nu = min_s + (int)(sintheta * (float)(max_u - min_s));
		const float inu = 1.f / (float)nu;
		rows[y].func.resize(nu);
		for(int x = 0; x < nu; ++x)
		{
			const float fx = ((float)x + 0.5f) * inu;
			const Col e = bgEval(bg, invSpheremap(fx, fy));
			rows[y].func[x] = ((e.r + e.g + e.b) * 0.333333f) * sintheta; // Rgb::energy color.h:58
		}
		cumulate(rows[y].func, nu, rows[y].cdf, rows[y].integral);
		fv[y] = rows[y].integral;
	}
	std::vector<float> vcdf; float vint;
	cumulate(fv, nv, vcdf, vint);
	d.cdf_offset = (int)tab.size(); d.cdf_count = nv; d.cdf_integral = vint;
	tab.insert(tab.end(), fv.begin(), fv.end());
	tab.insert(tab.end(), vcdf.begin(), vcdf.end());
	const size_t rowinfo = tab.size();
	tab.resize(rowinfo + 4 * (size_t)nv);
	for(int y = 0; y < nv; ++y)
	{
		const int off = (int)tab.size(), cnt = (int)rows[y].func.size();
		tab[rowinfo + 4 * y + 0] = i2f(off); tab[rowinfo + 4 * y + 1] = i2f(cnt);
		tab[rowinfo + 4 * y + 2] = rows[y].integral; tab[rowinfo + 4 * y + 3] = 1.f / rows[y].integral;
		tab.insert(tab.end(), rows[y].func.begin(), rows[y].func.end());
		tab.insert(tab.end(), rows[y].cdf.begin(), rows[y].cdf.end());
	}
}

// Camera::Camera camera.cc:46-66 + PerspectiveCamera ctor/setAxis camera_perspective.cc:29-82
inline bool prepCamera(const yafgpu_camera &c, CameraDev &d, std::string &err)
{
	memset(&d, 0, sizeof(d));
	if(c.type < YAFGPU_CAM_PERSPECTIVE || c.type > YAFGPU_CAM_EQUIRECTANGULAR) { err = "camera type not implemented on the device"; return false; }
	d.type = c.type;
	d.resx = c.resx; d.resy = c.resy;
	// Camera::Camera camera.cc:46-66
	const V3 pos = ld3(c.from), look = ld3(c.to), up = ld3(c.up);
	const float aspect_ratio = c.aspect_ratio * (float)c.resy / (float)c.resx;
	d.aspect_ratio = aspect_ratio;
	V3 cam_y = up - pos, cam_z = look - pos;
	V3 cam_x = cross(cam_z, cam_y);
	cam_y = cross(cam_z, cam_x);
	cam_x = normalize(cam_x); cam_y = normalize(cam_y); cam_z = normalize(cam_z);
	st3(d.near_n, cam_z);
	st3(d.near_p, pos + cam_z * c.near_clip);
	st3(d.far_p, pos + cam_z * c.far_clip);
	st3(d.pos, pos); st3(d.cam_to, cam_z);
	st3(d.axis_x, cam_x); st3(d.axis_y, cam_y); st3(d.axis_z, cam_z); d.focal = c.focal; // Camera::getAxis / screenproject (texture_mapper texco normal / window)
	if(c.type == YAFGPU_CAM_ORTHOGRAPHIC)
	{	// OrthographicCamera::setAxis camera_orthographic.cc:37-50
		V3 vright = cam_x, vup = aspect_ratio * cam_y;
		// pos_ = position_ - 0.5 * scale_ * (vup_ + vright_): (0.5 * scale_) is a double, operator*(float, Vec3) narrows it
		const V3 p0 = pos - (float)(0.5 * (double)c.scale) * (vup + vright);
		vup = vup * (c.scale / (float)c.resy);
		vright = vright * (c.scale / (float)c.resx);
		st3(d.pos, p0); st3(d.vright, vright); st3(d.vup, vup); d.ortho_scale = c.scale;
		return true;
	}
	if(c.type == YAFGPU_CAM_ANGULAR)
	{	// AngularCamera::AngularCamera camera_angular.cc:28-55, factory :100-133
		const float angle = (float)(c.angle * YG_M_PI / (double)180.f), max_angle = (float)(c.max_angle * YG_M_PI / (double)180.f);
		d.max_radius = max_angle / angle; d.circular = c.circular; d.projection = c.projection;
		if(c.projection == YAFGPU_AP_ORTHOGRAPHIC) d.focal_length = 1.f / fSin(angle);
		else if(c.projection == YAFGPU_AP_STEREOGRAPHIC) d.focal_length = (float)((double)(1.f / 2.f) / tan((double)(angle / 2.f)));
		else if(c.projection == YAFGPU_AP_EQUISOLID) d.focal_length = 1.f / 2.f / fSin(angle / 2.f);
		else if(c.projection == YAFGPU_AP_RECTILINEAR) d.focal_length = (float)((double)1.f / tan((double)angle));
		else d.focal_length = 1.f / angle;
		V3 vright = cam_x;
		if(c.mirrored) vright = vright * (float)-1.0;
		st3(d.vright, vright); st3(d.vup, cam_y);
		return true;
	}
	if(c.type == YAFGPU_CAM_EQUIRECTANGULAR)
	{	// EquirectangularCamera::setAxis camera_equirectangular.cc:38-48
		st3(d.vright, cam_x); st3(d.vup, cam_y);
		return true;
	}
	// PerspectiveCamera::PerspectiveCamera / setAxis camera_perspective.cc:29-74; ArchitectCamera::setAxis camera_architect.cc:52-66
	d.aperture = c.aperture; d.dof_distance = c.dof_distance;
	st3(d.dof_rt, c.aperture * cam_x); st3(d.dof_up, c.aperture * cam_y);
	V3 vright = cam_x;
	V3 vup = c.type == YAFGPU_CAM_ARCHITECT ? aspect_ratio * mk(0.f, 0.f, -1.f) : aspect_ratio * cam_y;
	// vto = (camZ * focal) - 0.5 * (vup + vright): 0.5 is double -> operator*(float, Vec3) narrows it to float first
	V3 vto = (cam_z * c.focal) - 0.5f * (vup + vright);
	vup = vup / (float)c.resy;
	vright = vright / (float)c.resx;
	st3(d.vright, vright); st3(d.vup, vup); st3(d.vto, vto);
	d.bokeh_type = c.bokeh_type; d.bokeh_bias = c.bokeh_bias; d.n_ls = 0;
	int ns = c.bokeh_type;
	if(ns >= 3 && ns <= 6)
	{
		float w = (float)((double)c.bokeh_rotation * 0.01745329251994329576922), wi = (float)(YG_M_2PI / (double)(float)ns);
		ns = (ns + 2) * 2;
		d.n_ls = ns;
		for(int i = 0; i < ns; i += 2) { d.ls[i] = fCos(w); d.ls[i + 1] = fSin(w); w += wi; }
	}
	return true;
}

// filter kernels imagefilm.cc:60-122
inline float filtBox(float, float) { return 1.f; }
inline float filtMitchell(float dx, float dy)
{
	float x = 2.f * fSqrt(dx * dx + dy * dy);
	if(x >= 2.f) return 0.f;
	if(x >= 1.f) return (float)(x * (x * (x * -0.38888889f + 2.0f) - 3.33333333f) + 1.77777778f);
	return (float)(x * x * (1.16666666f * x - 2.0f) + 0.88888889f);
}
inline float filtGauss(float dx, float dy)
{
	float r2 = dx * dx + dy * dy;
	// std::max(0.f, float(fExp__(-6 * r_2) - GAUSS_EXP)): GAUSS_EXP 0.00247875 is a double literal
	return fmaxf(0.f, (float)((double)fExp(-6 * r2) - 0.00247875));
}
inline float filtLanczos2(float dx, float dy)
{
	float x = fSqrt(dx * dx + dy * dy);
	if(x == 0.f) return 1.f;
	if(-2 < x && x < 2)
	{
		float a = (float)(YG_M_PI * (double)x);
		float b = (float)(1.57079632679489661923 * (double)x);
		return (fSin(a) * fSin(b)) / (a * b);
	}
	return 0.f;
}

// ImageFilm::ImageFilm imagefilm.cc:124-187
inline void prepFilm(const yafgpu_film_desc &f, FilmDev &d)
{
	d.w = f.width; d.h = f.height; d.cx0 = f.xstart; d.cy0 = f.ystart; d.cx1 = f.xstart + f.width; d.cy1 = f.ystart + f.height;
	d.filterw = (float)((double)f.aa_pixelwidth * 0.5);
	float (*fn)(float, float) = filtBox;
	switch(f.filter_type)
	{
		case YAFGPU_FILTER_MITCHELL: fn = filtMitchell; d.filterw *= 2.6f; break;
		case YAFGPU_FILTER_LANCZOS: fn = filtLanczos2; break;
		case YAFGPU_FILTER_GAUSS: fn = filtGauss; d.filterw *= 2.f; break;
		default: fn = filtBox;
	}
	d.filterw = fminf(fmaxf(0.501f, d.filterw), 0.5f * 8);
	const float scale = 1.f / 16.f;
	for(int y = 0; y < 16; ++y) for(int x = 0; x < 16; ++x) d.table[y * 16 + x] = fn((x + .5f) * scale, (y + .5f) * scale);
	d.table_scale = (float)(0.9999 * 16 / (double)d.filterw);
	d.clamp_samples = f.aa_clamp_samples; d.premult = f.premult; d.aa_passes = f.aa_passes;
	d.base_sampling_offset = f.base_sampling_offset;
}

} // namespace yg
