// GpuIntegrator — the reference-side adaptor of the yafgpu C ABI (include/yafgpu.h).
//
// This translation unit is compiled AGAINST THE REFERENCE'S HEADERS and linked into a libYafaRay front end in place of
// nothing: no reference source file is edited.  It plugs the GPU path in behind the reference's own integrator surface
//
//     class SurfaceIntegrator : Integrator        include/integrator/integrator.h:48-85
//         virtual bool preprocess();               called by Scene::update()            src/common/scene.cc:886
//         virtual bool render(int, ImageFilm *);   called by Scene::render()            src/common/scene.cc:1060
//         virtual void cleanup();                                                        src/common/scene.cc:1062
//
// so that scenes keep arriving through the unchanged XML loader / Interface API, and the reference keeps doing everything
// outside the hot path (parsing, smoothing, film flush, colour space, image output).
//
// Substitution happens at link time with GNU ld's --wrap (flags in oracle/Makefile, explained in INTEGRATION.md):
//   RenderEnvironment::create{Material,Light,Camera,Background,Integrator}, setupScene        src/common/environment.cc
// are wrapped to (a) keep a copy of the ParamMap each object was created from — the reference's factories consume the
// parameters and keep only constructor results — and (b) hand back a GpuIntegrator for `type` "pathtracing" and
// "directlighting" (what Integrator::factory would dispatch, src/integrator/integrator.cc:35-56).  A maintainer with
// commit access would instead add the two type strings to that factory and keep the ParamMaps in the environment tables.
//
// YAFGPU_DISABLE=1 in the environment leaves the reference's CPU integrators in place.  There is no CPU fallback: when the
// GPU path is selected and anything fails (no device, unsupported feature), preprocess()/render() log Y_ERROR and return false.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <limits>
#include <list>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <sstream>
#include "common/spectrum.h"
#include <string>
#include <thread>
#include <vector>

// read-only access to state the reference keeps private (mesh arrays, film buffers, environment tables); layouts unchanged
#define private public
#define protected public
#include "yafaray_config.h"
#include "common/environment.h"
#include "common/scene.h"
#include "common/imagefilm.h"
#include "common/param.h"
#include "common/session.h"
#include "common/logging.h"
#include "common/timer.h"
#include "common/triangle.h"
#include "common/monitor.h"
#include "object_geom/object_geom_mesh.h"
#include "material/material.h"
#include "material/material_node.h"
#include "light/light.h"
#include "camera/camera.h"
#include "background/background.h"
#include "background/background_texture.h"
#include "texture/texture_image.h"
#include "texture/texture_basic.h"
#include "imagehandler/imagehandler.h"
#include "integrator/integrator.h"
#undef private
#undef protected

#include "yafgpu.h"

BEGIN_YAFARAY

// the Voronoi distance metrics have external linkage but no declaration in the reference's headers (noise_generator.cc:377-409)
float distRealF__(float, float, float, float); float distSquaredF__(float, float, float, float); float distChebychevF__(float, float, float, float);
float distMinkovskyHf__(float, float, float, float); float distMinkovsky4F__(float, float, float, float); float distMinkovskyF__(float, float, float, float);


namespace {

// ---- what the wrapped factories remember ----------------------------------------------------------------------------
struct Recorded
{
	std::map<const void *, ParamMap> params;     // object -> the ParamMap it was created from
	std::map<const void *, std::string> names;
	std::vector<const Material *> materials;     // creation order = material index on the device
	std::map<const void *, std::list<ParamMap>> eparams; // material -> its <list_element> maps (shader nodes)
	RenderEnvironment *env = nullptr;
	ParamMap film;                               // the <render> block as setupScene saw it
	ParamMap integrator;
	std::string background_name;
	std::list<std::string> ies_texts;            // contents of the IES files of the scene's ieslights (kept alive across yafgpu_upload_scene)
};
Recorded &rec() { static Recorded r; return r; }

template <class T> T par(const ParamMap &p, const char *key, T def) { p.getParam(key, def); return def; }
void par3(const ParamMap &p, const char *key, float *out, float d0, float d1, float d2)
{
	Point3 v(d0, d1, d2);
	p.getParam(key, v);
	out[0] = v.x_; out[1] = v.y_; out[2] = v.z_;
}
void parCol(const ParamMap &p, const char *key, float *out, float d)
{
	Rgb c(d);
	p.getParam(key, c);
	out[0] = c.r_; out[1] = c.g_; out[2] = c.b_;
}
bool hasShaderNodes(const ParamMap &p)
{
	for(auto it = p.begin(); it != p.end(); ++it)
	{
		const std::string &k = it->first;
		if(k.size() > 7 && k.compare(k.size() - 7, 7, "_shader") == 0) return true;
	}
	return false;
}

// ---- ParamMap -> C ABI records (same keys and in-code defaults as the reference factories) ------------------------------
// Scene-wide node and texture tables handed to yafgpu_upload_scene (include/yafgpu.h: yafgpu_shader_node, yafgpu_texture)
struct NodePool
{
	std::vector<yafgpu_shader_node> nodes;
	std::vector<yafgpu_texture> textures;
	std::map<const Texture *, int> tex_index;
	std::list<std::vector<float>> texels; // stable storage behind yafgpu_texture::rgba
};

// ImageTexture -> yafgpu_texture: parameters from the constructed object (the factory's results, texture_image.cc:560-720), texels
// from the reference's own ImageHandler::getPixel — whatever format the reference could load, colour-space decode included
bool textureRecord(const Texture *t, NodePool &pool, int &index, std::string &err)
{
	auto f = pool.tex_index.find(t);
	if(f != pool.tex_index.end()) { index = f->second; return true; }
	yafgpu_texture r;
	memset(&r, 0, sizeof(r));
	auto noiseBasis = [&](const NoiseGenerator *ng, int &out) -> bool
	{	// the basis is recognised by the dynamic type newNoise__ (texture_basic.cc:24-52) constructed
		if(dynamic_cast<const StdPerlinNoiseGenerator *>(ng)) out = YAFGPU_NOISE_STDPERLIN;
		else if(dynamic_cast<const BlenderNoiseGenerator *>(ng)) out = YAFGPU_NOISE_BLENDER;
		else if(dynamic_cast<const CellNoiseGenerator *>(ng)) out = YAFGPU_NOISE_CELL;
		else if(const VoronoiNoiseGenerator *vg = dynamic_cast<const VoronoiNoiseGenerator *>(ng))
		{
			if(vg->distfunc_2_ != distRealF__) { err = "a Voronoi noise basis with a distance metric other than the factory's default is not implemented on the device"; return false; }
			out = YAFGPU_NOISE_VORONOI_F1 + (int)vg->v_type_;   // Vf1, Vf2, Vf3, Vf4, Vf2F1, VCrackle: the enumerator order of YAFGPU_NOISE_VORONOI_*
		}
		else if(dynamic_cast<const NewPerlinNoiseGenerator *>(ng)) out = YAFGPU_NOISE_NEWPERLIN;
		else { err = "unknown noise basis"; return false; }
		return true;
	};
	auto adjustments = [&]()
	{
		r.adj_intensity = t->adj_intensity_; r.adj_contrast = t->adj_contrast_; r.adj_saturation = t->adj_saturation_; r.adj_hue = t->adj_hue_;
		r.adj_mult_red = t->adj_mult_factor_red_; r.adj_mult_green = t->adj_mult_factor_green_; r.adj_mult_blue = t->adj_mult_factor_blue_;
		r.adj_clamp = t->adj_clamp_;
	};
	auto done = [&]() { index = (int)pool.textures.size(); pool.textures.push_back(r); pool.tex_index[t] = index; return true; };
	auto colorRamp = [&]() -> bool
	{	// ColorRamp (include/common/color_ramp.h): the items as addItem keeps them (sorted by position), the enums in declaration order
		const ColorRamp *cr = t->color_ramp_;
		if(!cr) return true;
		pool.texels.emplace_back();
		std::vector<float> &v = pool.texels.back();
		for(const ColorRampItem &it : cr->ramp_) { const Rgba c = it.color(); v.push_back(c.r_); v.push_back(c.g_); v.push_back(c.b_); v.push_back(c.a_); v.push_back(it.position()); }
		r.ramp_items = (int)cr->ramp_.size(); r.ramp = v.data();
		r.ramp_mode = (int)cr->mode_;                                      // Rgb, Hsv, Hsl
		r.ramp_interpolation = cr->interpolation_ == ColorRamp::Constant ? 0 : 1;
		r.ramp_hue_interpolation = cr->hue_interpolation_ == ColorRamp::Far ? YAFGPU_RAMPHUE_FAR : cr->hue_interpolation_ == ColorRamp::Clockwise ? YAFGPU_RAMPHUE_CW :
		                           cr->hue_interpolation_ == ColorRamp::Counterclockwise ? YAFGPU_RAMPHUE_CCW : YAFGPU_RAMPHUE_NEAR;
		return true;
	};
	const BlendTexture *bt = dynamic_cast<const BlendTexture *>(t);
	const RgbCubeTexture *ct = dynamic_cast<const RgbCubeTexture *>(t);
	if(bt || ct)
	{	// texture_basic.cc:680-779, 347-386: fields of the constructed objects
		if(bt && !colorRamp()) return false;
		r.type = bt ? YAFGPU_TEX_BLEND : YAFGPU_TEX_RGB_CUBE;
		if(bt) { r.blend_type = (int)bt->progression_type_; r.use_flip_axis = bt->use_flip_axis_; } // same enumerator order as YAFGPU_TEXBLEND_*
		r.adj_intensity = t->adj_intensity_; r.adj_contrast = t->adj_contrast_; r.adj_saturation = t->adj_saturation_; r.adj_hue = t->adj_hue_;
		r.adj_mult_red = t->adj_mult_factor_red_; r.adj_mult_green = t->adj_mult_factor_green_; r.adj_mult_blue = t->adj_mult_factor_blue_;
		r.adj_clamp = t->adj_clamp_;
		index = (int)pool.textures.size();
		pool.textures.push_back(r);
		pool.tex_index[t] = index;
		return true;
	}
	const VoronoiTexture *vot = dynamic_cast<const VoronoiTexture *>(t);
	const MusgraveTexture *mut = dynamic_cast<const MusgraveTexture *>(t);
	const DistortedNoiseTexture *dnt = dynamic_cast<const DistortedNoiseTexture *>(t);
	if(vot || mut || dnt)
	{	// texture_basic.cc:391-673: the members of the constructed objects
		if(!colorRamp()) return false;
		if(vot)
		{
			r.type = YAFGPU_TEX_VORONOI; r.noise_size = vot->size_; r.color_type = vot->coltype_; r.intensity_scale = vot->iscale_; r.mk_exponent = vot->v_gen_.mk_exp_;
			r.vor_weight[0] = vot->w_1_; r.vor_weight[1] = vot->w_2_; r.vor_weight[2] = vot->w_3_; r.vor_weight[3] = vot->w_4_;
			// setDistM (noise_generator.cc:487-521) only swaps the function pointer (and maps "manhattan" to the squared metric): read the metric off it
			const DistMetricFunc_t fn = vot->v_gen_.distfunc_2_;
			r.distance_metric = fn == distSquaredF__ ? YAFGPU_DIST_SQUARED : fn == distChebychevF__ ? YAFGPU_DIST_CHEBYCHEV : fn == distMinkovskyHf__ ? YAFGPU_DIST_MINKOVSKY_HALF :
			                    fn == distMinkovsky4F__ ? YAFGPU_DIST_MINKOVSKY_FOUR : fn == distMinkovskyF__ ? YAFGPU_DIST_MINKOVSKY : YAFGPU_DIST_REAL;
		}
		else if(mut)
		{
			r.type = YAFGPU_TEX_MUSGRAVE; r.noise_size = mut->size_; r.intensity_scale = mut->iscale_;
			if(!noiseBasis(mut->n_gen_, r.noise_type)) return false;
			r.color1[0] = mut->color_1_.r_; r.color1[1] = mut->color_1_.g_; r.color1[2] = mut->color_1_.b_;
			r.color2[0] = mut->color_2_.r_; r.color2[1] = mut->color_2_.g_; r.color2[2] = mut->color_2_.b_;
			r.mus_offset = 1.f; r.mus_gain = 1.f;
			if(const MFractalMusgrave *m = dynamic_cast<const MFractalMusgrave *>(mut->m_gen_)) { r.musgrave_type = YAFGPU_MUSGRAVE_MULTIFRACTAL; r.mus_h = m->h_; r.mus_lacunarity = m->lacunarity_; r.mus_octaves = m->octaves_; }
			else if(const HeteroTerrainMusgrave *m = dynamic_cast<const HeteroTerrainMusgrave *>(mut->m_gen_)) { r.musgrave_type = YAFGPU_MUSGRAVE_HETEROTERRAIN; r.mus_h = m->h_; r.mus_lacunarity = m->lacunarity_; r.mus_octaves = m->octaves_; r.mus_offset = m->offset_; }
			else if(const HybridMFractalMusgrave *m = dynamic_cast<const HybridMFractalMusgrave *>(mut->m_gen_)) { r.musgrave_type = YAFGPU_MUSGRAVE_HYBRIDMF; r.mus_h = m->h_; r.mus_lacunarity = m->lacunarity_; r.mus_octaves = m->octaves_; r.mus_offset = m->offset_; r.mus_gain = m->gain_; }
			else if(const RidgedMFractalMusgrave *m = dynamic_cast<const RidgedMFractalMusgrave *>(mut->m_gen_)) { r.musgrave_type = YAFGPU_MUSGRAVE_RIDGEDMF; r.mus_h = m->h_; r.mus_lacunarity = m->lacunarity_; r.mus_octaves = m->octaves_; r.mus_offset = m->offset_; r.mus_gain = m->gain_; }
			else if(const FBmMusgrave *m = dynamic_cast<const FBmMusgrave *>(mut->m_gen_)) { r.musgrave_type = YAFGPU_MUSGRAVE_FBM; r.mus_h = m->h_; r.mus_lacunarity = m->lacunarity_; r.mus_octaves = m->octaves_; }
			else { err = "unknown Musgrave generator"; return false; }
		}
		else
		{
			r.type = YAFGPU_TEX_DISTORTED_NOISE; r.noise_size = dnt->size_; r.distort = dnt->distort_;
			if(!noiseBasis(dnt->n_gen_1_, r.noise_type) || !noiseBasis(dnt->n_gen_2_, r.noise_type2)) return false;
			r.color1[0] = dnt->color_1_.r_; r.color1[1] = dnt->color_1_.g_; r.color1[2] = dnt->color_1_.b_;
			r.color2[0] = dnt->color_2_.r_; r.color2[1] = dnt->color_2_.g_; r.color2[2] = dnt->color_2_.b_;
		}
		adjustments();
		return done();
	}
	const CloudsTexture *clt = dynamic_cast<const CloudsTexture *>(t);
	const MarbleTexture *mbt = dynamic_cast<const MarbleTexture *>(t);
	const WoodTexture *wdt = dynamic_cast<const WoodTexture *>(t);
	if(clt || mbt || wdt)
	{	// texture_basic.cc:96-343: the members the constructors left behind; the basis is recognised by the dynamic type of n_gen_ (newNoise__ :24-52)
		if(!colorRamp()) return false;
		if(!noiseBasis(clt ? clt->n_gen_ : mbt ? mbt->n_gen_ : wdt->n_gen_, r.noise_type)) return false;
		const Rgb c1 = clt ? clt->color_1_ : mbt ? mbt->color_1_ : wdt->color_1_, c2 = clt ? clt->color_2_ : mbt ? mbt->color_2_ : wdt->color_2_;
		r.color1[0] = c1.r_; r.color1[1] = c1.g_; r.color1[2] = c1.b_; r.color2[0] = c2.r_; r.color2[1] = c2.g_; r.color2[2] = c2.b_;
		r.sharpness = 1.f;
		if(clt) { r.type = YAFGPU_TEX_CLOUDS; r.noise_depth = clt->depth_; r.noise_size = clt->size_; r.noise_hard = clt->hard_; r.clouds_bias = clt->bias_; }
		else if(mbt)
		{
			r.type = YAFGPU_TEX_MARBLE; r.noise_depth = mbt->octaves_; r.noise_size = mbt->size_; r.noise_hard = mbt->hard_;
			r.turbulence = mbt->turb_; r.sharpness = mbt->sharpness_; r.wave_shape = (int)mbt->wshape_;
		}
		else
		{
			r.type = YAFGPU_TEX_WOOD; r.noise_depth = wdt->octaves_; r.noise_size = wdt->size_; r.noise_hard = wdt->hard_;
			r.turbulence = wdt->turb_; r.wave_shape = (int)wdt->wshape_; r.wood_rings = wdt->rings_;
		}
		adjustments();
		return done();
	}
	const ImageTexture *it = dynamic_cast<const ImageTexture *>(t);
	if(!it) { err = "this procedural texture type is not implemented on the device (image, blend, rgb_cube, clouds, marble, wood, voronoi, musgrave and distorted_noise are)"; return false; }
	r.type = YAFGPU_TEX_IMAGE;
	ImageHandler *ih = it->image_;
	r.width = ih->getWidth(); r.height = ih->getHeight();
	pool.texels.emplace_back((size_t)r.width * r.height * 4);
	std::vector<float> &px = pool.texels.back();
	for(int y = 0; y < r.height; ++y) for(int x = 0; x < r.width; ++x)
	{
		const Rgba c = ih->getPixel(x, y);
		float *o = &px[((size_t)y * r.width + x) * 4];
		o[0] = c.r_; o[1] = c.g_; o[2] = c.b_; o[3] = c.a_;
	}
	r.rgba = px.data();
	switch(it->getInterpolationType())
	{
		case InterpolationType::None: r.interpolate = YAFGPU_TEXINTERP_NONE; break;
		case InterpolationType::Bilinear: r.interpolate = YAFGPU_TEXINTERP_BILINEAR; break;
		case InterpolationType::Bicubic: r.interpolate = YAFGPU_TEXINTERP_BICUBIC; break;
		default: err = "texture interpolation mipmap_trilinear / mipmap_ewa (ray differentials) is not implemented on the device"; return false;
	}
	r.clip_mode = (int)it->tex_clip_mode_; // same enumerator order as YAFGPU_TEXCLIP_*
	r.xrepeat = it->xrepeat_; r.yrepeat = it->yrepeat_; r.mirror_x = it->mirror_x_; r.mirror_y = it->mirror_y_; r.rot90 = it->rot_90_;
	r.cropmin_x = it->cropminx_; r.cropmin_y = it->cropminy_; r.cropmax_x = it->cropmaxx_; r.cropmax_y = it->cropmaxy_;
	r.checker_even = it->checker_even_; r.checker_odd = it->checker_odd_; r.checker_dist = it->checker_dist_;
	r.normalmap = it->normalmap_; r.color_space = (int)it->color_space_; r.gamma = it->gamma_;
	r.adj_intensity = it->adj_intensity_; r.adj_contrast = it->adj_contrast_; r.adj_saturation = it->adj_saturation_;
	r.adj_hue = it->adj_hue_;        // already "adj_hue" / 60.f (setAdjustments, texture.h:154), which is what the ABI field holds
	r.adj_mult_red = it->adj_mult_factor_red_; r.adj_mult_green = it->adj_mult_factor_green_; r.adj_mult_blue = it->adj_mult_factor_blue_;
	r.adj_clamp = it->adj_clamp_;
	index = (int)pool.textures.size();
	pool.textures.push_back(r);
	pool.tex_index[t] = index;
	return true;
}

// NodeMaterial::loadNodes / parseNodes (material_node.cc:140-237) restated on the recorded <list_element> maps: the nodes in declaration
// order and the "<slot>_shader" bindings.  When the reference itself ended up without nodes (loadNodes failed, e.g. an image the build
// has no decoder for) the material has none here either.
bool nodeRecords(const Material *m, const ParamMap &p, const std::string &name, yafgpu_material &out, NodePool &pool, std::string &err)
{
	for(int s = 0; s < YAFGPU_N_SLOTS; ++s) out.shader[s] = -1;
	out.node_first = (int32_t)pool.nodes.size(); out.node_count = 0;
	const NodeMaterial *nm = dynamic_cast<const NodeMaterial *>(m);
	if(!nm || nm->m_shaders_table_.empty()) return true;
	auto ep = rec().eparams.find(m);
	if(ep == rec().eparams.end()) return true;
	std::vector<const ParamMap *> elems;
	for(const ParamMap &e : ep->second)
	{
		std::string element;
		if(e.getParam("element", element) && element != "shader_node") continue;
		elems.push_back(&e);
	}
	std::map<std::string, int> idx;
	for(size_t i = 0; i < elems.size(); ++i) idx[par<std::string>(*elems[i], "name", "")] = (int)i;
	auto ref = [&](const ParamMap &e, const char *key) -> int
	{
		std::string n;
		if(!e.getParam(key, n)) return -1;
		auto f = idx.find(n);
		return f == idx.end() ? -1 : f->second;
	};
	auto col4 = [&](const ParamMap &e, const char *key, float *o)
	{
		Rgba c(0.f, 0.f, 0.f, 1.f);
		e.getParam(key, c);
		o[0] = c.r_; o[1] = c.g_; o[2] = c.b_; o[3] = c.a_;
	};
	for(const ParamMap *ee : elems)
	{
		const ParamMap &e = *ee;
		const std::string type = par<std::string>(e, "type", "");
		yafgpu_shader_node r;
		memset(&r, 0, sizeof(r));
		r.input1 = r.input2 = r.factor = r.texture = -1;
		if(type == "texture_mapper")
		{	// shader_node_basic.cc:344-420
			r.type = YAFGPU_NODE_TEXTURE_MAPPER;
			const Texture *tex = rec().env ? rec().env->getTexture(par<std::string>(e, "texture", "")) : nullptr;
			if(!tex) { err = "material '" + name + "': texture of a texture_mapper node not found"; return false; }
			int ti = -1;
			if(!textureRecord(tex, pool, ti, err)) { err = "material '" + name + "': " + err; return false; }
			r.texture = ti;
			const std::string tc = par<std::string>(e, "texco", "global"), mp = par<std::string>(e, "mapping", "plain");
			r.texco = tc == "uv" ? YAFGPU_TEXCO_UV : tc == "orco" ? YAFGPU_TEXCO_ORCO : tc == "transformed" ? YAFGPU_TEXCO_TRANSFORMED : tc == "window" ? YAFGPU_TEXCO_WINDOW
			        : tc == "normal" ? YAFGPU_TEXCO_NORMAL : tc == "reflect" ? YAFGPU_TEXCO_REFLECT : tc == "stick" ? YAFGPU_TEXCO_STICK : tc == "stress" ? YAFGPU_TEXCO_STRESS
			        : tc == "tangent" ? YAFGPU_TEXCO_TANGENT : YAFGPU_TEXCO_GLOBAL;
			r.mapping = mp == "cube" ? YAFGPU_TEXPROJ_CUBE : mp == "tube" ? YAFGPU_TEXPROJ_TUBE : mp == "sphere" ? YAFGPU_TEXPROJ_SPHERE : YAFGPU_TEXPROJ_PLAIN;
			r.proj_x = par<int>(e, "proj_x", 1); r.proj_y = par<int>(e, "proj_y", 2); r.proj_z = par<int>(e, "proj_z", 3);
			par3(e, "scale", r.scale, 1.f, 1.f, 1.f); par3(e, "offset", r.offset, 0.f, 0.f, 0.f);
			r.bump_strength = par<float>(e, "bump_strength", 1.f); r.do_scalar = par<bool>(e, "do_scalar", true);
			Matrix4 mtx(1);
			e.getParam("transform", mtx);
			for(int a = 0; a < 4; ++a) for(int b = 0; b < 4; ++b) r.transform[4 * a + b] = mtx[a][b];
		}
		else if(type == "value")
		{
			r.type = YAFGPU_NODE_VALUE;
			Rgb c(1.f);
			e.getParam("color", c);
			r.color[0] = c.r_; r.color[1] = c.g_; r.color[2] = c.b_; r.color[3] = par<float>(e, "alpha", 1.f);
			r.scalar = par<float>(e, "scalar", 1.f);
		}
		else if(type == "mix")
		{
			r.type = YAFGPU_NODE_MIX; r.mode = par<int>(e, "mode", 0);
			r.input1 = ref(e, "input1"); r.input2 = ref(e, "input2"); r.factor = ref(e, "factor");
			col4(e, "color1", r.color); col4(e, "color2", r.color2);
			r.scalar = par<float>(e, "value", 0.f);
		}
		else if(type == "layer")
		{	// shader_node_layer.cc:160-244
			r.type = YAFGPU_NODE_LAYER; r.mode = par<int>(e, "mode", 0);
			r.input1 = ref(e, "input"); r.input2 = ref(e, "upper_layer");
			Rgb dc(1.f);
			e.getParam("def_col", dc);
			r.color[0] = dc.r_; r.color[1] = dc.g_; r.color[2] = dc.b_; r.color[3] = 1.f;
			col4(e, "upper_color", r.color2);
			r.scalar = (float)par<double>(e, "def_val", 1.0); r.scalar2 = par<float>(e, "upper_value", 0.f);
			r.colfac = (float)par<double>(e, "colfac", 1.0); r.valfac = (float)par<double>(e, "valfac", 1.0);
			r.do_color = par<bool>(e, "do_color", true); r.layer_do_scalar = par<bool>(e, "do_scalar", false);
			r.color_input = par<bool>(e, "color_input", true); r.use_alpha = par<bool>(e, "use_alpha", false);
			r.no_rgb = par<bool>(e, "noRGB", false); r.stencil = par<bool>(e, "stencil", false); r.negative = par<bool>(e, "negative", false);
		}
		else { err = "material '" + name + "': shader node type '" + type + "' is not implemented on the device"; return false; }
		pool.nodes.push_back(r);
	}
	out.node_count = (int32_t)elems.size();
	static const char *slot_keys[YAFGPU_N_SLOTS] = {"diffuse_shader", "mirror_color_shader", "bump_shader", "mirror_shader", "transparency_shader", "translucency_shader",
		"sigma_oren_shader", "diffuse_refl_shader", "IOR_shader", "wireframe_shader", "glossy_shader", "glossy_reflect_shader", "exponent_shader", "filter_color_shader",
		"roughness_shader", "mask"};
	const std::string mtype = par<std::string>(p, "type", "");
	for(int s = 0; s < YAFGPU_N_SLOTS; ++s)
	{
		if(s == YAFGPU_SLOT_BLEND && mtype != "blend_mat" && mtype != "mask_mat") continue;
		out.shader[s] = ref(p, slot_keys[s]);
	}
	return true;
}

bool materialRecord(const Material *m, yafgpu_material &out, NodePool &pool, std::string &err)
{
	memset(&out, 0, sizeof(out));
	auto it = rec().params.find(m);
	if(it == rec().params.end()) { err = "material was not created through RenderEnvironment::createMaterial"; return false; }
	const ParamMap &p = it->second;
	const std::string name = rec().names[m];
	if(!nodeRecords(m, p, name, out, pool, err)) return false;
	const std::string type = par<std::string>(p, "type", "");
	const std::string vis = par<std::string>(p, "visibility", "normal");     // material factories: e.g. material_shiny_diffuse.cc:626-649
	out.visibility = vis == "no_shadows" ? YAFGPU_VIS_NO_SHADOWS : vis == "shadow_only" ? YAFGPU_VIS_SHADOW_ONLY : vis == "invisible" ? YAFGPU_VIS_INVISIBLE : YAFGPU_VIS_NORMAL;
	out.receive_shadows = par<bool>(p, "receive_shadows", true);
	out.flat_material = par<bool>(p, "flat_material", false);
	out.additional_depth = par<int>(p, "additionaldepth", 0);
	out.transparent_bias_factor = par<float>(p, "transparentbias_factor", 0.f);
	out.transparent_bias_multiply_raydepth = par<bool>(p, "transparentbias_multiply_raydepth", false);
	out.sampling_factor = par<float>(p, "samplingfactor", 1.f);
	out.mat_pass_index = par<int>(p, "mat_pass_index", 0);
	auto indexOf = [&](const std::string &child) -> int
	{
		for(size_t i = 0; i < rec().materials.size(); ++i) if(rec().names[rec().materials[i]] == child) return (int)i;
		return -1;
	};
	if(type == "shinydiffusemat")
	{	// material_shiny_diffuse.cc:599-688
		out.type = YAFGPU_MAT_SHINYDIFFUSE;
		yafgpu_mat_shinydiffuse &s = out.u.shinydiffuse;
		parCol(p, "color", s.color, 1.f); parCol(p, "mirror_color", s.mirror_color, 1.f);
		s.diffuse_reflect = par<float>(p, "diffuse_reflect", 1.f); s.specular_reflect = par<float>(p, "specular_reflect", 0.f);
		s.transparency = par<float>(p, "transparency", 0.f); s.translucency = par<float>(p, "translucency", 0.f);
		s.emit = par<float>(p, "emit", 0.f); s.ior = par<float>(p, "IOR", 1.33f);
		s.fresnel_effect = par<bool>(p, "fresnel_effect", false); s.transmit_filter = par<float>(p, "transmit_filter", 1.f);
		s.oren_nayar = par<std::string>(p, "diffuse_brdf", "") == "oren_nayar"; s.sigma = par<float>(p, "sigma", 0.1f);
	}
	else if(type == "glass" || type == "rough_glass")
	{	// material_glass.cc:311-379, material_rough_glass.cc:322-401
		out.type = type == "glass" ? YAFGPU_MAT_GLASS : YAFGPU_MAT_ROUGH_GLASS;
		yafgpu_mat_glass &s = out.u.glass;
		s.ior = par<float>(p, "IOR", 1.4f);
		parCol(p, "filter_color", s.filter_color, 1.f); parCol(p, "mirror_color", s.mirror_color, 1.f);
		s.transmit_filter = par<float>(p, "transmit_filter", 0.f); s.dispersion_power = par<float>(p, "dispersion_power", 0.f);
		s.fake_shadows = par<bool>(p, "fake_shadows", false); s.alpha = par<float>(p, "alpha", 0.5f);
		Rgb absorb(1.f);
		const bool has = p.getParam("absorption", absorb);
		s.has_absorption = has && (absorb.r_ < 1.f || absorb.g_ < 1.f || absorb.b_ < 1.f) && m->vol_i_ != nullptr; // the factory made a Beer handler (needs "name")
		s.absorption[0] = absorb.r_; s.absorption[1] = absorb.g_; s.absorption[2] = absorb.b_;
		s.absorption_dist = par<float>(p, "absorption_dist", 1.f);
	}
	else if(type == "glossy" || type == "coated_glossy")
	{	// material_glossy.cc:407-480, material_coated_glossy.cc:465-545
		out.type = type == "glossy" ? YAFGPU_MAT_GLOSSY : YAFGPU_MAT_COATED_GLOSSY;
		yafgpu_mat_glossy &s = out.u.glossy;
		parCol(p, "color", s.color, 1.f); parCol(p, "diffuse_color", s.diffuse_color, 1.f); parCol(p, "mirror_color", s.mirror_color, 1.f);
		s.diffuse_reflect = par<float>(p, "diffuse_reflect", 0.f); s.glossy_reflect = par<float>(p, "glossy_reflect", 1.f);
		s.specular_reflect = par<float>(p, "specular_reflect", 1.f);
		s.exponent = par<float>(p, "exponent", 50.f); s.exp_u = par<float>(p, "exp_u", 50.f); s.exp_v = par<float>(p, "exp_v", 50.f);
		s.ior = par<float>(p, "IOR", 1.4f);
		s.as_diffuse = par<bool>(p, "as_diffuse", true); s.anisotropic = par<bool>(p, "anisotropic", false);
		s.oren_nayar = par<std::string>(p, "diffuse_brdf", "") == "Oren-Nayar"; s.sigma = par<float>(p, "sigma", 0.1f);
	}
	else if(type == "mask_mat")
	{	// material_mask.cc:150-198
		out.type = YAFGPU_MAT_MASK;
		// the mask factory reads nothing but threshold, the two children, receive_shadows and visibility: the rest stays at Material's defaults
		out.flat_material = 0; out.additional_depth = 0; out.transparent_bias_factor = 0.f; out.transparent_bias_multiply_raydepth = 0;
		out.sampling_factor = 1.f; out.mat_pass_index = 0;
		out.u.blend.material1 = indexOf(par<std::string>(p, "material1", ""));
		out.u.blend.material2 = indexOf(par<std::string>(p, "material2", ""));
		out.u.blend.blend_value = (float)par<double>(p, "threshold", 0.5);
	}
	else if(type == "blend_mat")
	{	// material_blend.cc:464-518
		out.type = YAFGPU_MAT_BLEND;
		out.u.blend.material1 = indexOf(par<std::string>(p, "material1", ""));
		out.u.blend.material2 = indexOf(par<std::string>(p, "material2", ""));
		out.u.blend.blend_value = par<float>(p, "blend_value", 0.5f);
	}
	else if(type == "light_mat")
	{	// material_simple.cc:63-73
		out.type = YAFGPU_MAT_LIGHT;
		parCol(p, "color", out.u.simple.color, 1.f);
		out.u.simple.strength = par<float>(p, "power", 1.f);
		out.u.simple.double_sided = par<bool>(p, "double_sided", false);
	}
	else if(type == "null") out.type = YAFGPU_MAT_NULL;
	else { err = "material '" + name + "': type '" + type + "' is not implemented on the device"; return false; }
	return true;
}

bool lightRecord(const Light *l, const std::map<int, int> &obj_index_of_id, yafgpu_light &out, std::string &err)
{
	memset(&out, 0, sizeof(out));
	auto it = rec().params.find(l);
	if(it == rec().params.end()) { err = "light was not created through RenderEnvironment::createLight"; return false; }
	const ParamMap &p = it->second;
	const std::string type = par<std::string>(p, "type", "");
	out.light_enabled = par<bool>(p, "light_enabled", true); out.cast_shadows = par<bool>(p, "cast_shadows", true);
	out.photon_only = par<bool>(p, "photon_only", false);
	out.shoot_caustic = par<bool>(p, "with_caustic", true); out.shoot_diffuse = par<bool>(p, "with_diffuse", true);
	parCol(p, "color", out.color, 1.f);
	out.power = par<float>(p, "power", 1.f);
	out.samples = par<int>(p, "samples", 4);
	out.object = -1;
	auto objectOf = [&]() -> int
	{
		const int id = par<int>(p, "object", 0);
		auto f = obj_index_of_id.find(id);
		return (id != 0 && f != obj_index_of_id.end()) ? f->second : -1;
	};
	if(type == "pointlight") { out.type = YAFGPU_LIGHT_POINT; par3(p, "from", out.from, 0.f, 0.f, 0.f); }               // light_point.cc:101-133
	else if(type == "sunlight") { out.type = YAFGPU_LIGHT_SUN; par3(p, "direction", out.from, 0.f, 0.f, 1.f); out.angle = par<float>(p, "angle", 0.27f); } // light_sun.cc:95-125
	else if(type == "arealight")
	{	// light_area.cc:174-213
		out.type = YAFGPU_LIGHT_AREA;
		par3(p, "corner", out.from, 0.f, 0.f, 0.f); par3(p, "point1", out.point1, 0.f, 0.f, 0.f); par3(p, "point2", out.point2, 0.f, 0.f, 0.f);
		out.object = objectOf();
	}
	else if(type == "bgPortalLight")
	{	// light_background_portal.cc:236-266
		out.type = YAFGPU_LIGHT_BGPORTAL;
		out.object = objectOf();
	}
	else if(type == "meshlight")
	{	// light_meshlight.cc:228-263
		out.type = YAFGPU_LIGHT_MESH;
		out.object = objectOf();
		out.double_sided = par<bool>(p, "double_sided", false);
	}
	else if(type == "bglight")
	{	// light_background.cc:251-283 (created by a background with ibl=true, background_constant.cc:69-83)
		out.type = YAFGPU_LIGHT_BACKGROUND;
		out.samples = par<int>(p, "samples", 16);
		out.abs_intersect = par<bool>(p, "abs_intersect", false);
		out.clamp_intersect = l->clamp_intersect_; // set after creation by TextureBackground::factory (background_texture.cc:155-160)
	}
	else if(type == "directional")
	{	// light_directional.cc:112-158
		out.type = YAFGPU_LIGHT_DIRECTIONAL;
		par3(p, "direction", out.point1, 0.f, 0.f, 1.f);
		out.infinite = par<bool>(p, "infinite", true);
		if(!out.infinite)
		{
			Point3 from(0.f);
			if(!p.getParam("from", from)) p.getParam("position", from);
			out.from[0] = from.x_; out.from[1] = from.y_; out.from[2] = from.z_;
			out.radius = par<float>(p, "radius", 1.f);
		}
	}
	else if(type == "spotlight")
	{	// light_spot.cc:251-295
		out.type = YAFGPU_LIGHT_SPOT;
		par3(p, "from", out.from, 0.f, 0.f, 0.f); par3(p, "to", out.point1, 0.f, 0.f, -1.f);
		out.cone_angle = par<float>(p, "cone_angle", 45.f); out.blend = par<float>(p, "blend", 0.15f);
		out.soft_shadows = par<bool>(p, "soft_shadows", false); out.shadow_fuzzy = par<float>(p, "shadowFuzzyness", 1.f);
		out.samples = par<int>(p, "samples", 8);
	}
	else if(type == "ieslight")
	{	// light_ies.cc:184-233; the library parses the file contents as IesData::parseIesFile does (the light exists, so the reference's parse succeeded)
		out.type = YAFGPU_LIGHT_IES;
		par3(p, "from", out.from, 0.f, 0.f, 0.f); par3(p, "to", out.point1, 0.f, 0.f, -1.f);
		out.soft_shadows = par<bool>(p, "soft_shadows", false); out.samples = par<int>(p, "samples", 16);
		const std::string file = par<std::string>(p, "file", "");
		std::ifstream fin(file.c_str(), std::ios::in | std::ios::binary);
		if(!fin) { err = "cannot read the IES file '" + file + "'"; return false; }
		std::stringstream ss; ss << fin.rdbuf();
		rec().ies_texts.push_back(ss.str());
		out.ies_data = rec().ies_texts.back().data(); out.ies_size = rec().ies_texts.back().size();
	}
	else if(type == "spherelight")
	{	// light_sphere.cc:163-198
		out.type = YAFGPU_LIGHT_SPHERE;
		par3(p, "from", out.from, 0.f, 0.f, 0.f); out.radius = par<float>(p, "radius", 1.f);
		out.object = par<int>(p, "object", 0); // a linked sphere primitive is refused by the library
	}
	else { err = "light type '" + type + "' is not implemented on the device"; return false; }
	return true;
}

} // namespace

// ---- the integrator ---------------------------------------------------------------------------------------------------
class GpuIntegrator final : public SurfaceIntegrator
{
	public:
		explicit GpuIntegrator(const ParamMap &params) : params_(params)
		{
			type_ = Surface;
			integrator_name_ = "GPU " + par<std::string>(params, "type", "");
			integrator_short_name_ = "GPU";
		}
		~GpuIntegrator() override { if(grp_) yafgpu_group_destroy(grp_); }
		bool preprocess() override;
		bool render(int num_view, ImageFilm *image_film) override;
		void cleanup() override {}
		Rgba integrate(RenderState &, DiffRay &, ColorPasses &, int) const override { return Rgba(0.f); } // never called: render() is overridden

	private:
		bool fail(const std::string &what)
		{
			Y_ERROR << integrator_name_ << ": " << what << (grp_ ? std::string(" — ") + yafgpu_group_last_error(grp_) : std::string()) << YENDL;
			return false;
		}
		bool fillIntegrator(yafgpu_integrator_desc &it, std::string &err) const;
		bool depositFilm(ImageFilm *film);
		ParamMap params_;
		yafgpu_group *grp_ = nullptr;   // the GPUs that render this scene: YAFGPU_DEVICES="0,1,2,3" | "all", else YAFGPU_DEVICE (one GPU, default 0)
		std::vector<float> host_film_, bg_texels_;
};

bool GpuIntegrator::fillIntegrator(yafgpu_integrator_desc &it, std::string &err) const
{
	memset(&it, 0, sizeof(it));
	const ParamMap &p = params_;
	const std::string type = par<std::string>(p, "type", "");
	it.raydepth = par<int>(p, "raydepth", 5); it.shadow_depth = par<int>(p, "shadowDepth", 5);
	it.transp_shad = par<bool>(p, "transpShad", false);
	it.bg_transp = par<bool>(p, "bg_transp", false); it.bg_transp_refract = par<bool>(p, "bg_transp_refract", false);
	it.do_ao = par<bool>(p, "do_AO", false); it.ao_samples = par<int>(p, "AO_samples", 32); it.ao_distance = par<float>(p, "AO_distance", 1.f);
	parCol(p, "AO_color", it.ao_color, 1.f);
	if(type == "pathtracing")
	{	// integrator_path_tracer.cc:349-423
		it.type = YAFGPU_INT_PATHTRACING;
		it.path_samples = par<int>(p, "path_samples", 32); it.bounces = par<int>(p, "bounces", 3);
		it.rr_min_bounces = par<int>(p, "russian_roulette_min_bounces", 0); it.no_recursive = par<bool>(p, "no_recursive", false);
		const std::string c = par<std::string>(p, "caustic_type", "path");
		if(c == "photon" || c == "both") { err = "caustic photon maps are not implemented on the device"; return false; }
		it.trace_caustics = c != "none";
	}
	else if(type == "directlighting")
	{	// integrator_direct_light.cc:196-250
		it.type = YAFGPU_INT_DIRECTLIGHT;
		if(par<bool>(p, "caustics", false)) { err = "caustic photon maps are not implemented on the device"; return false; }
	}
	else { err = "integrator type '" + type + "'"; return false; }
	it.shadow_bias = scene_->shadow_bias_; it.shadow_bias_auto = scene_->shadow_bias_auto_;
	it.ray_min_dist = scene_->ray_min_dist_; it.ray_min_dist_auto = scene_->ray_min_dist_auto_;
	return true;
}

// Replaces the accelerator-building half of Scene::update() (scene.cc:788-831) for this integrator: flatten what the
// reference has built (meshes after smoothing, materials, lights, camera, background, film settings) and upload it.
bool GpuIntegrator::preprocess()
{
	std::string err;
	if(!grp_)
	{	// device list: integrator parameter "gpu_devices" (string) or YAFGPU_DEVICES, "all" = every GPU of the box; else one GPU (YAFGPU_DEVICE, default 0)
		std::string list = par<std::string>(params_, "gpu_devices", "");
		if(const char *env = getenv("YAFGPU_DEVICES")) list = env;
		std::vector<int> devs;
		bool all = list == "all";
		if(!all && !list.empty())
		{
			std::stringstream ss(list);
			std::string tok;
			while(std::getline(ss, tok, ',')) if(!tok.empty()) devs.push_back(atoi(tok.c_str()));
		}
		if(!all && devs.empty()) { const char *dev = getenv("YAFGPU_DEVICE"); devs.push_back(dev ? atoi(dev) : 0); }
		if(yafgpu_group_create(&grp_, all ? nullptr : devs.data(), all ? 0 : (int)devs.size(), 0) != YAFGPU_OK)
			return fail(std::string("cannot create the CUDA contexts: ") + yafgpu_group_last_error(nullptr));
	}
	yafgpu_scene_desc d;
	memset(&d, 0, sizeof(d));
	d.abi_version = YAFGPU_ABI_VERSION;

	// materials in creation order
	std::vector<yafgpu_material> mats(std::max<size_t>(1, rec().materials.size()));
	std::map<const Material *, int> mat_index;
	NodePool pool;
	for(size_t i = 0; i < rec().materials.size(); ++i)
	{
		if(!materialRecord(rec().materials[i], mats[i], pool, err)) return fail(err);
		mat_index[rec().materials[i]] = (int)i;
	}

	// geometry in the reference's global triangle order: ascending ObjId_t, visible non-base triangle meshes (scene.cc:796-817)
	std::vector<float> pts, nrm, uvs;
	std::vector<int32_t> vidx, nidx, uvidx, tmat, tobj;
	std::vector<yafgpu_object> objs;
	std::map<int, int> obj_index_of_id;
	bool any_n = false, any_uv = false;
	// BackgroundPortalLight::init hides its mesh (light_background_portal.cc:90), but only after Scene::update has put it into the kd-tree this
	// render uses (scene.cc:788-831 run before the light loop): the portal meshes are part of the geometry
	std::set<int> portal_objects;
	for(const Light *l : scene_->lights_)
	{
		auto it = rec().params.find(l);
		if(it != rec().params.end() && par<std::string>(it->second, "type", "") == "bgPortalLight") portal_objects.insert(par<int>(it->second, "object", 0));
	}
	for(auto &kv : scene_->meshes_)
	{
		ObjData &dat = kv.second;
		if(dat.type_ != 0 /* TRIM */) return fail("only triangle meshes (type 0) are implemented on the device");
		const TriangleObject *o = dat.obj_;
		if((!o->isVisible() && !portal_objects.count((int)kv.first)) || o->isBaseObject()) continue;
		const int p_off = (int)(pts.size() / 3), n_off = (int)(nrm.size() / 3), uv_off = (int)(uvs.size() / 2);
		yafgpu_object rec_o;
		memset(&rec_o, 0, sizeof(rec_o));
		rec_o.has_orco = o->has_orco_; rec_o.has_uv = o->has_uv_; rec_o.smooth = o->is_smooth_ || o->normals_exported_; rec_o.light = -1;
		rec_o.first_tri = (int32_t)tmat.size(); rec_o.n_tris = (int32_t)o->triangles_.size(); rec_o.obj_id = (int32_t)(kv.first & 0x7fffffff);
		obj_index_of_id[(int)kv.first] = (int)objs.size();
		for(const Point3 &p : o->points_) { pts.push_back(p.x_); pts.push_back(p.y_); pts.push_back(p.z_); }
		for(const Normal &n : o->normals_) { nrm.push_back(n.x_); nrm.push_back(n.y_); nrm.push_back(n.z_); }
		for(const Uv &uv : o->uv_values_) { uvs.push_back(uv.u_); uvs.push_back(uv.v_); }
		for(const Triangle &t : o->triangles_)
		{
			vidx.push_back(t.pa_ + p_off); vidx.push_back(t.pb_ + p_off); vidx.push_back(t.pc_ + p_off);
			const int na[3] = {t.na_, t.nb_, t.nc_};
			for(int k = 0; k < 3; ++k) { const bool has = rec_o.smooth && na[k] >= 0; nidx.push_back(has ? na[k] + n_off : -1); any_n |= has; }
			for(int k = 0; k < 3; ++k)
			{
				if(o->has_uv_) { uvidx.push_back(o->uv_offsets_[t.self_index_ * 3 + k] + uv_off); any_uv = true; }
				else uvidx.push_back(-1);
			}
			auto mi = mat_index.find(t.material_);
			if(mi == mat_index.end()) return fail("a triangle uses a material that was not created through the environment");
			tmat.push_back(mi->second);
			tobj.push_back((int32_t)objs.size());
		}
		objs.push_back(rec_o);
	}
	yafgpu_geometry &g = d.geometry;
	g.points = pts.data(); g.n_points = pts.size() / 3;
	g.normals = nrm.data(); g.n_normals = nrm.size() / 3;
	g.uv = uvs.data(); g.n_uv = uvs.size() / 2;
	g.tri_vidx = vidx.data(); g.tri_nidx = any_n ? nidx.data() : nullptr; g.tri_uvidx = any_uv ? uvidx.data() : nullptr;
	g.tri_mat = tmat.data(); g.tri_obj = tobj.data(); g.n_tris = tmat.size();
	g.objects = objs.data(); g.n_objects = objs.size();
	d.materials = mats.data(); d.n_materials = rec().materials.size();
	d.nodes = pool.nodes.data(); d.n_nodes = pool.nodes.size();
	d.textures = pool.textures.data(); d.n_textures = pool.textures.size();

	// lights in Scene::lights_ order (the order doLightEstimation's per-light QMC offsets depend on)
	rec().ies_texts.clear();
	std::vector<yafgpu_light> lights(std::max<size_t>(1, scene_->lights_.size()));
	for(size_t i = 0; i < scene_->lights_.size(); ++i)
		if(!lightRecord(scene_->lights_[i], obj_index_of_id, lights[i], err)) return fail(err);
	d.lights = lights.data(); d.n_lights = scene_->lights_.size();

	// camera factories: camera_perspective.cc:206-252, camera_architect.cc:100-133, camera_orthographic.cc:82-106,
	// camera_angular.cc:100-133, camera_equirectangular.cc:82-108
	{
		auto it = rec().params.find(scene_->getCamera());
		if(it == rec().params.end()) return fail("camera was not created through RenderEnvironment::createCamera");
		const ParamMap &p = it->second;
		const std::string ctype = par<std::string>(p, "type", "");
		yafgpu_camera &c = d.camera;
		if(ctype == "perspective") c.type = YAFGPU_CAM_PERSPECTIVE;
		else if(ctype == "architect") c.type = YAFGPU_CAM_ARCHITECT;
		else if(ctype == "orthographic") c.type = YAFGPU_CAM_ORTHOGRAPHIC;
		else if(ctype == "angular") c.type = YAFGPU_CAM_ANGULAR;
		else if(ctype == "equirectangular") c.type = YAFGPU_CAM_EQUIRECTANGULAR;
		else return fail("camera type '" + ctype + "' is not implemented on the device");
		c.resx = par<int>(p, "resx", 320); c.resy = par<int>(p, "resy", 200);
		par3(p, "from", c.from, 0.f, 1.f, 0.f); par3(p, "to", c.to, 0.f, 0.f, 0.f); par3(p, "up", c.up, 0.f, 1.f, 1.f);
		c.focal = par<float>(p, "focal", 1.f); c.aperture = par<float>(p, "aperture", 0.f); c.dof_distance = par<float>(p, "dof_distance", 0.f);
		c.aspect_ratio = par<float>(p, "aspect_ratio", 1.f); c.near_clip = par<float>(p, "nearClip", 0.f);
		c.far_clip = par<float>(p, "farClip", (ctype == "angular" || ctype == "equirectangular") ? -1.0e38f : -1.f);
		c.scale = par<float>(p, "scale", 1.f);
		c.angle = par<double>(p, "angle", 90.0); c.max_angle = par<double>(p, "max_angle", c.angle);
		c.circular = par<bool>(p, "circular", true); c.mirrored = par<bool>(p, "mirrored", false);
		const std::string proj = par<std::string>(p, "projection", "");
		c.projection = proj == "orthographic" ? YAFGPU_AP_ORTHOGRAPHIC : proj == "stereographic" ? YAFGPU_AP_STEREOGRAPHIC
		             : proj == "equisolid_angle" ? YAFGPU_AP_EQUISOLID : proj == "rectilinear" ? YAFGPU_AP_RECTILINEAR : YAFGPU_AP_EQUIDISTANT;
		const std::string bt = par<std::string>(p, "bokeh_type", "disk1"), bb = par<std::string>(p, "bokeh_bias", "uniform");
		c.bokeh_type = bt == "disk2" ? YAFGPU_BK_DISK2 : bt == "triangle" ? YAFGPU_BK_TRI : bt == "square" ? YAFGPU_BK_SQR : bt == "pentagon" ? YAFGPU_BK_PENTA
		             : bt == "hexagon" ? YAFGPU_BK_HEXA : bt == "ring" ? YAFGPU_BK_RING : YAFGPU_BK_DISK1;
		c.bokeh_bias = bb == "center" ? YAFGPU_BB_CENTER : bb == "edge" ? YAFGPU_BB_EDGE : YAFGPU_BB_NONE;
		c.bokeh_rotation = par<float>(p, "bokeh_rotation", 0.f);
	}

	// background (background_constant.cc:52-86)
	d.background.type = YAFGPU_BG_NONE;
	if(const Background *bg = scene_->getBackground())
	{
		auto it = rec().params.find(bg);
		if(it == rec().params.end()) return fail("background was not created through RenderEnvironment::createBackground");
		const ParamMap &p = it->second;
		const std::string btype = par<std::string>(p, "type", "");
		if(btype == "constant")
		{
			d.background.type = YAFGPU_BG_CONSTANT;
			parCol(p, "color", d.background.color, 0.f);
		}
		else if(btype == "textureback")
		{	// background_texture.cc:90-165; texels are taken from the reference's own ImageHandler (post load, post colour-space handling)
			const TextureBackground *tb = static_cast<const TextureBackground *>(bg);
			const ImageTexture *it_ = dynamic_cast<const ImageTexture *>(tb->tex_);
			if(!it_) return fail("textureback: only image textures are implemented on the device");
			if(it_->tex_clip_mode_ != ImageTexture::TexClipMode::Repeat || it_->xrepeat_ != 1 || it_->yrepeat_ != 1 || it_->mirror_x_ || it_->mirror_y_
			   || it_->cropx_ || it_->cropy_ || it_->rot_90_ || it_->adjustments_set_)
				return fail("textureback: texture clipping/repeat/mirror/crop/rot90/adjustments are not implemented on the device");
			const InterpolationType ip = it_->getInterpolationType();
			if(ip != InterpolationType::Bilinear && ip != InterpolationType::None) return fail("textureback: only 'bilinear' and 'none' texture interpolation are implemented on the device");
			if(par<float>(p, "smartibl_blur", 0.f) > 0.f) return fail("textureback: smartibl_blur is not implemented on the device");
			ImageHandler *ih = it_->image_;
			const int tw = ih->getWidth(), th = ih->getHeight();
			bg_texels_.resize((size_t)tw * th * 3);
			for(int y = 0; y < th; ++y) for(int x = 0; x < tw; ++x)
			{
				const Rgba c = ih->getPixel(x, y);
				float *o = &bg_texels_[((size_t)y * tw + x) * 3];
				o[0] = c.r_; o[1] = c.g_; o[2] = c.b_;
			}
			d.background.type = YAFGPU_BG_TEXTURE;
			d.background.tex_rgb = bg_texels_.data(); d.background.tex_w = tw; d.background.tex_h = th;
			d.background.tex_interp = ip == InterpolationType::Bilinear;
			const std::string mapping = par<std::string>(p, "mapping", "");
			d.background.mapping_angular = mapping == "probe" || mapping == "angular";
			d.background.rotation = par<float>(p, "rotation", 0.f);
		}
		else if(btype == "gradientback")
		{	// background_gradient.cc:66-117
			yafgpu_background &b = d.background;
			b.type = YAFGPU_BG_GRADIENT;
			parCol(p, "horizon_color", b.horizon_color, 1.f);
			{ Rgb z(0.4f, 0.5f, 1.f); p.getParam("zenith_color", z); b.zenith_color[0] = z.r_; b.zenith_color[1] = z.g_; b.zenith_color[2] = z.b_; }
			{ Rgb c(b.horizon_color[0], b.horizon_color[1], b.horizon_color[2]); p.getParam("horizon_ground_color", c); b.horizon_ground_color[0] = c.r_; b.horizon_ground_color[1] = c.g_; b.horizon_ground_color[2] = c.b_; }
			{ Rgb c(b.zenith_color[0], b.zenith_color[1], b.zenith_color[2]); p.getParam("zenith_ground_color", c); b.zenith_ground_color[0] = c.r_; b.zenith_ground_color[1] = c.g_; b.zenith_ground_color[2] = c.b_; }
		}
		else return fail("background type '" + btype + "' is not implemented on the device");
		d.background.power = par<float>(p, "power", 1.f); d.background.ibl = par<bool>(p, "ibl", false);
		d.background.with_caustic = par<bool>(p, "with_caustic", true);
	}

	// film + AA (environment.cc:456-584, 679-813); the ImageFilm the reference created is the authority for the geometry
	{
		const ParamMap &p = rec().film;
		const ImageFilm *film = scene_->getImageFilm();
		yafgpu_film_desc &f = d.film;
		f.width = film->w_; f.height = film->h_; f.xstart = film->cx_0_; f.ystart = film->cy_0_;
		f.aa_pixelwidth = par<float>(p, "AA_pixelwidth", 1.5f);
		const std::string ft = par<std::string>(p, "filter_type", "box");
		f.filter_type = ft == "mitchell" ? YAFGPU_FILTER_MITCHELL : ft == "gauss" ? YAFGPU_FILTER_GAUSS : ft == "lanczos" ? YAFGPU_FILTER_LANCZOS : YAFGPU_FILTER_BOX;
		f.tile_size = film->getTileSize(); f.premult = par<bool>(p, "premult", false);
		f.aa_clamp_samples = par<float>(p, "AA_clamp_samples", 0.f); f.aa_clamp_indirect = par<float>(p, "AA_clamp_indirect", 0.f);
		f.aa_passes = par<int>(p, "AA_passes", 1);
		f.aa_light_sample_multiplier = f.aa_indirect_sample_multiplier = 1.f;
		f.base_sampling_offset = film->getBaseSamplingOffset();
	}
	if(!fillIntegrator(d.integrator, err)) return fail(err);
	// CIE table for dispersive glass: rows 360..829 nm are wl2Xyz__ at integer wavelengths (the interpolation weights are 1 and 0 there), row 830
	// follows from the midpoint 829.5 nm; only 400..700 nm is ever looked up (spectrum.h:41-46)
	std::vector<float> cie(471 * 3, 0.f);
	for(int i = 0; i < 470; ++i) { const Rgb c = wl2Xyz__(360.f + (float)i); cie[3 * i] = c.r_; cie[3 * i + 1] = c.g_; cie[3 * i + 2] = c.b_; }
	{ const Rgb c = wl2Xyz__(829.5f); cie[3 * 470] = 2.f * c.r_ - cie[3 * 469]; cie[3 * 470 + 1] = 2.f * c.g_ - cie[3 * 469 + 1]; cie[3 * 470 + 2] = 2.f * c.b_ - cie[3 * 469 + 2]; }
	d.cie_xyz = cie.data(); d.n_cie_rows = 471;
	// lattice tables of the noise bases: file-static data of src/texture/noise_generator.cc, compiled in from the header that
	// oracle/ref_driver/noise_dump.cc writes out of that very source when the integration is built (oracle/Makefile `gpu`)
	static const unsigned char noise_tables[] = {
#include "noise_tables.inc"
	};
	static_assert(sizeof(noise_tables) == YAFGPU_NOISE_TABLES_BYTES, "noise table blob");
	d.noise_tables = noise_tables; d.noise_tables_bytes = sizeof(noise_tables);
	if(yafgpu_group_upload_scene(grp_, &d) != YAFGPU_OK) return fail("scene upload refused");
	yafgpu_stats st;
	yafgpu_group_get_stats(grp_, &st);
	Y_INFO << integrator_name_ << ": " << g.n_tris << " triangles on " << yafgpu_group_size(grp_) << " GPU(s), BVH built on the GPU in " << st.ms_build << " ms" << YENDL;
	return true;
}

// GPU accumulators -> the reference's film buffers (Pixel{Rgba col_; float weight_}, util_image_buffers.h:36-47), Combined pass
bool GpuIntegrator::depositFilm(ImageFilm *film)
{
	const int w = film->w_, h = film->h_;
	host_film_.resize((size_t)w * h * 5);
	if(yafgpu_group_read_film(grp_, host_film_.data()) != YAFGPU_OK) return fail("film read-back failed");
	Rgba2DImageWeighed_t &img = *film->image_passes_[0];
	for(int y = 0; y < h; ++y) for(int x = 0; x < w; ++x)
	{
		const float *s = &host_film_[((size_t)y * w + x) * 5];
		Pixel &p = img(x, y);
		p.col_ = Rgba(s[0], s[1], s[2], s[3]);
		p.weight_ = s[4];
	}
	return true;
}

// The pass schedule of TiledIntegrator::render (integrator_tiled.cc:116-259) with renderPass replaced by yafgpu_render_pass.
// Adaptive resampling uses the reference's own ImageFilm::nextPass on the deposited film to flag pixels (imagefilm.cc:270-480).
bool GpuIntegrator::render(int num_view, ImageFilm *film)
{
	int aa_samples, aa_passes, aa_inc; float thr, floor_pct, smf, lmf, imf, dark_factor, clamp_s, clamp_i; bool color_noise; DarkDetectionType dark_type; int var_edge, var_pix;
	scene_->getAaParameters(aa_samples, aa_passes, aa_inc, thr, floor_pct, smf, lmf, imf, color_noise, dark_type, dark_factor, var_edge, var_pix, clamp_s, clamp_i);
	session__.setStatusTotalPasses(aa_passes);
	float sample_mult = 1.f, light_mult = 1.f, indirect_mult = 1.f;
	const int floor_pixels = (int)floorf(floor_pct * (float)film->getTotalPixels() / 100.f);
	g_timer__.addEvent("rendert");
	g_timer__.start("rendert");
	film->init(aa_passes);
	film->setAaNoiseParams(color_noise, dark_type, dark_factor, var_edge, var_pix, clamp_s);
	const int w = film->w_, h = film->h_;
	const unsigned base = film->getBaseSamplingOffset();
	if(session__.renderResumed())
	{	// film_save_load "load-save": ImageFilm::init has summed the YAF_FILMv1 files of the output folder into the film (imagefilm.cc:1467-1558):
		// the device continues from those accumulators
		host_film_.resize((size_t)w * h * 5);
		const Rgba2DImageWeighed_t &img = *film->image_passes_[0];
		for(int y = 0; y < h; ++y) for(int x = 0; x < w; ++x)
		{
			const Pixel &p = img(x, y);
			float *o = &host_film_[((size_t)y * w + x) * 5];
			o[0] = p.col_.r_; o[1] = p.col_.g_; o[2] = p.col_.b_; o[3] = p.col_.a_; o[4] = p.weight_;
		}
		if(yafgpu_group_write_film(grp_, host_film_.data()) != YAFGPU_OK) return fail("film upload failed");
	}
	else if(yafgpu_group_film_clear(grp_) != YAFGPU_OK) return fail("film clear failed");

	auto pass = [&](int samples, int offset, bool adaptive, int number, const uint8_t *mask) -> bool
	{
		yafgpu_pass_desc pd;
		memset(&pd, 0, sizeof(pd));
		pd.n_samples = samples; pd.offset = (int32_t)(offset + base); pd.adaptive = adaptive; pd.aa_pass_number = number; pd.resample_mask = mask;
		session__.setStatusCurrentPass(number + 1);
		film->setSamplingOffset(offset + samples);
		if(yafgpu_group_render_pass(grp_, &pd) != YAFGPU_OK) return fail("render pass failed");
		return depositFilm(film);
	};
	int acum = aa_samples, resampled = 0;
	if(session__.renderResumed())
	{	// "skipping pass 1": renderPass(0 samples, offset = the loaded film's sampling offset) (integrator_tiled.cc:198-202)
		acum = (int)film->getSamplingOffset();
		session__.setStatusCurrentPass(1);
		film->setSamplingOffset(acum);
	}
	else if(!pass(aa_samples, 0, false, 0, nullptr)) return false;
	bool thr_changed = true;
	std::vector<uint8_t> mask;
	for(int i = 1; i < aa_passes; ++i)
	{
		if(scene_->getSignals() & Y_SIG_ABORT) break;
		sample_mult *= smf; light_mult *= lmf; indirect_mult *= imf;
		yafgpu_group_set_aa_multipliers(grp_, light_mult, indirect_mult);
		film->setAaNoiseParams(color_noise, dark_type, dark_factor, var_edge, var_pix, clamp_s);
		if(resampled <= 0 && !thr_changed) film->nextPass(num_view, true, integrator_name_, /*skip_next_pass=*/true);
		else
		{
			film->setAaThreshold(thr);
			resampled = film->nextPass(num_view, true, integrator_name_);
			thr_changed = false;
		}
		const int n = (int)ceilf(aa_inc * sample_mult);
		if(resampled > 0)
		{
			const uint8_t *m = nullptr;
			if(thr > 0.f)
			{	// ImageFilm::doMoreSamples (imagefilm.cc:917-920)
				mask.assign((size_t)w * h, 0);
				for(int y = 0; y < h; ++y) for(int x = 0; x < w; ++x) mask[(size_t)y * w + x] = film->doMoreSamples(x + film->cx_0_, y + film->cy_0_) ? 1 : 0;
				m = mask.data();
			}
			if(!pass(n, acum, true, i, m)) return false;
		}
		acum += n;
		if(resampled < floor_pixels)
		{
			const float ratio = std::min(8.f, (float)floor_pixels / resampled);
			thr *= (1.f - 0.1f * ratio);
			if(thr > 0.f) thr_changed = true;
		}
	}
	g_timer__.stop("rendert");
	session__.setStatusRenderFinished();
	yafgpu_stats st;
	yafgpu_group_get_stats(grp_, &st);
	const double rays = (double)(st.rays_camera + st.rays_extension + st.rays_shadow + st.rays_shadow_transp);
	Y_INFO << integrator_name_ << ": Overall rendertime: " << g_timer__.getTime("rendert") << "s, " << rays / 1e6 << " Mrays, "
	       << (st.ms_total > 0 ? rays / st.ms_total / 1e3 : 0.0) << " Mrays/s on the device" << YENDL;
	return true;
}

END_YAFARAY

// ---- link-time interception (GNU ld --wrap=<mangled name>, see oracle/Makefile `GPU_WRAP`) ----------------------------------
// `__wrap_X` receives every call to X made from another object file; `__real_X` is the reference's own definition.
// Member functions are declared as free functions taking `this` first (Itanium C++ ABI: identical calling convention).
using namespace yafaray4;
#define YG_STR "NSt7__cxx1112basic_stringIcSt11char_traitsIcESaIcEEE"
#define YG_ENV "_ZN8yafaray417RenderEnvironment"
#define YG_SYM_MATERIAL   YG_ENV "14createMaterialERK" YG_STR "RNS_8ParamMapERNS1_4listIS9_SaIS9_EEE"
#define YG_SYM_LIGHT      YG_ENV "11createLightERK" YG_STR "RNS_8ParamMapE"
#define YG_SYM_CAMERA     YG_ENV "12createCameraERK" YG_STR "RNS_8ParamMapE"
#define YG_SYM_BACKGROUND YG_ENV "16createBackgroundERK" YG_STR "RNS_8ParamMapE"
#define YG_SYM_INTEGRATOR YG_ENV "16createIntegratorERK" YG_STR "RNS_8ParamMapE"
#define YG_SYM_SETUP      YG_ENV "10setupSceneERNS_5SceneERKNS_8ParamMapERNS_11ColorOutputEPNS_11ProgressBarE"

extern "C" {
Material *yg_real_material(RenderEnvironment *, const std::string &, ParamMap &, std::list<ParamMap> &) asm("__real_" YG_SYM_MATERIAL);
Light *yg_real_light(RenderEnvironment *, const std::string &, ParamMap &) asm("__real_" YG_SYM_LIGHT);
Camera *yg_real_camera(RenderEnvironment *, const std::string &, ParamMap &) asm("__real_" YG_SYM_CAMERA);
Background *yg_real_background(RenderEnvironment *, const std::string &, ParamMap &) asm("__real_" YG_SYM_BACKGROUND);
Integrator *yg_real_integrator(RenderEnvironment *, const std::string &, ParamMap &) asm("__real_" YG_SYM_INTEGRATOR);
bool yg_real_setup(RenderEnvironment *, Scene &, const ParamMap &, ColorOutput &, ProgressBar *) asm("__real_" YG_SYM_SETUP);

Material *yg_wrap_material(RenderEnvironment *, const std::string &, ParamMap &, std::list<ParamMap> &) asm("__wrap_" YG_SYM_MATERIAL);
Light *yg_wrap_light(RenderEnvironment *, const std::string &, ParamMap &) asm("__wrap_" YG_SYM_LIGHT);
Camera *yg_wrap_camera(RenderEnvironment *, const std::string &, ParamMap &) asm("__wrap_" YG_SYM_CAMERA);
Background *yg_wrap_background(RenderEnvironment *, const std::string &, ParamMap &) asm("__wrap_" YG_SYM_BACKGROUND);
Integrator *yg_wrap_integrator(RenderEnvironment *, const std::string &, ParamMap &) asm("__wrap_" YG_SYM_INTEGRATOR);
bool yg_wrap_setup(RenderEnvironment *, Scene &, const ParamMap &, ColorOutput &, ProgressBar *) asm("__wrap_" YG_SYM_SETUP);
}

static void remember(const void *obj, const std::string &name, const ParamMap &p)
{
	if(!obj) return;
	rec().params[obj] = p;
	rec().names[obj] = name;
}

Material *yg_wrap_material(RenderEnvironment *env, const std::string &name, ParamMap &p, std::list<ParamMap> &e)
{
	const ParamMap copy = p; // factories may edit the map they are given
	Material *m = yg_real_material(env, name, p, e);
	if(m) { remember(m, name, copy); rec().materials.push_back(m); rec().eparams[m] = e; rec().env = env; }
	return m;
}
Light *yg_wrap_light(RenderEnvironment *env, const std::string &name, ParamMap &p)
{
	const ParamMap copy = p;
	Light *l = yg_real_light(env, name, p);
	remember(l, name, copy);
	return l;
}
Camera *yg_wrap_camera(RenderEnvironment *env, const std::string &name, ParamMap &p)
{
	const ParamMap copy = p;
	Camera *c = yg_real_camera(env, name, p);
	remember(c, name, copy);
	return c;
}
Background *yg_wrap_background(RenderEnvironment *env, const std::string &name, ParamMap &p)
{
	const ParamMap copy = p;
	Background *b = yg_real_background(env, name, p);
	remember(b, name, copy);
	return b;
}
// setupScene receives the <render> block (film size, filter, AA settings): environment.cc:679-813
bool yg_wrap_setup(RenderEnvironment *env, Scene &scene, const ParamMap &p, ColorOutput &out, ProgressBar *pb)
{
	rec().film = p;
	return yg_real_setup(env, scene, p, out, pb);
}
// What Integrator::factory would do with two more `else if` lines (src/integrator/integrator.cc:35-56).
Integrator *yg_wrap_integrator(RenderEnvironment *env, const std::string &name, ParamMap &p)
{
	std::string type;
	p.getParam("type", type);
	const char *off = getenv("YAFGPU_DISABLE");
	bool gpu = !(off && atoi(off) != 0);
	p.getParam("gpu", gpu);
	if(!gpu || (type != "pathtracing" && type != "directlighting")) return yg_real_integrator(env, name, p);
	if(env->integrators_.find(name) != env->integrators_.end()) return nullptr; // same rule as createIntegrator (environment.cc:370-373)
	Integrator *i = new GpuIntegrator(p);
	env->integrators_[name] = i;
	Y_INFO << "Environment: GPU integrator substituted for '" << type << "' (" << name << ")" << YENDL;
	return i;
}
