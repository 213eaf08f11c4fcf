// TEST INFRASTRUCTURE — not product code.  Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline /
// --impl reference legs may execute the binaries built from this file.
//
// Front end that links the UNMODIFIED reference sources (oracle/Makefile) and exposes the reference's own CPU
// implementation of the hot path for parity checks and CPU timing:
//
//   yaf_ref render <scene.xml> <out.film> [threads]   Scene::render() through the XML loader exactly as
//                                                     /root/reference/src/loader_xml/loader_xml.cc:217-308 does, then dumps
//                                                     the raw ImageFilm accumulators (Pixel{Rgba col_, float weight_},
//                                                     util_image_buffers.h:36-47) of the Combined pass.
//   yaf_ref hits <scene.xml> <rays.bin> <hits.bin> [mode] [shadow_depth]
//                                                     Scene::intersect / Scene::isShadowed (scene.cc:896-1035) per ray,
//                                                     global triangle ids in the order of scene.cc:807-817.
//   yaf_ref bench <scene.xml> <threads> <warmup> <steps> [spp]
//                                                     CPU baseline of bench.py: loads the scene ONCE (XML parse and Scene::update, i.e. the kd-tree
//                                                     build, timed separately), then calls Scene::render() warmup + steps times; every call is one
//                                                     full-frame render at `spp` samples per pixel (AA_minsamples override) and reports the
//                                                     reference's own "rendert" timer.
//   yaf_ref kat <out.bin>                             known-answer vectors of the QMC / fast-math / filter helpers.
//
// File formats (little endian): rays.bin = n x {from[3], tmin, dir[3], tmax} float32 (yafgpu_ray);
// hits.bin = n x {t,u,v float32, tri int32} (yafgpu_hit); out.film = "YGF1", int32 w, h, then w*h*5 float32.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <limits>
#include <list>
#include <map>
#include <mutex>
#include <set>
#include <sstream>
#include <string>
#include <thread>
#include <vector>
#include <condition_variable>
#include <iomanip>
#include <memory>

// Open access to the reference's private state for read-only inspection (object layout is unchanged).
#define private public
#define protected public
#include "yafaray_config.h"
#include "common/environment.h"
#include "common/scene.h"
#include "common/imagefilm.h"
#include "common/import_xml.h"
#include "common/param.h"
#include "common/session.h"
#include "common/logging.h"
#include "common/timer.h"
#include "common/surface.h"
#include "common/triangle.h"
#include "common/kdtree_triangle.h"
#include "common/scr_halton.h"
#include "common/spectrum.h"
#include "object_geom/object_geom_mesh.h"
#include "material/material.h"
#include "light/light.h"
#include "background/background.h"
#include "camera/camera.h"
#include "integrator/integrator.h"
#include "output/output.h"
#include "utility/util_mcqmc.h"
#include "utility/util_sample.h"
#include "utility/util_math_optimizations.h"
#undef private
#undef protected

using namespace yafaray4;

namespace {

// A sink that drops everything: the film accumulators are read directly instead (MemoryInputOutput is unusable,
// /root/reference/src/common/output_memory.cc:39-41 writes r, g and b to one channel).
class NullOutput : public ColorOutput
{
	public:
		bool putPixel(int, int, int, const RenderPasses *, int, const Rgba &, bool) override { return true; }
		bool putPixel(int, int, int, const RenderPasses *, const std::vector<Rgba> &, bool) override { return true; }
		void flush(int, const RenderPasses *) override {}
		void flushArea(int, int, int, int, int, const RenderPasses *) override {}
		// ImageFilm::flush saves the "YAF_FILMv1" file only for an image output (imagefilm.cc:900-906); YAF_REF_FILM_PATH plays the role of
		// yafaray-xml's output path (output_image.cc:36), so `film_save_load` in the <render> block works as it does there
		bool isImageOutput() override { return getenv("YAF_REF_FILM_PATH") != nullptr; }
};

struct Loaded
{
	RenderEnvironment *env = nullptr;
	Scene *scene = nullptr;
	NullOutput out;
	ParamMap render;
};

bool loadScene(const char *xml, Loaded &l, int threads)
{
	logger__.setConsoleMasterVerbosity(getenv("YAF_REF_VERBOSE") ? "verbose" : "warning");
	logger__.setLogMasterVerbosity("mute");
	l.env = new RenderEnvironment();
	l.scene = new Scene(l.env);
	l.env->setScene(l.scene);
	if(!parseXmlFile__(xml, l.scene, l.env, l.render, "LinearRGB", 1.f)) return false;
	if(threads >= -1) l.render["threads"] = threads;
	l.render["logging_saveLog"] = false;
	l.render["logging_saveHTML"] = false;
	if(const char *fp = getenv("YAF_REF_FILM_PATH")) session__.setPathImageOutput(fp);
	if(!l.env->setupScene(*l.scene, l.render, l.out)) return false;
	session__.setInteractive(false);
	return true;
}

int cmdRender(int argc, char **argv)
{
	if(argc < 4) { fprintf(stderr, "usage: yaf_ref render scene.xml out.film [threads]\n"); return 2; }
	int threads = argc > 4 ? atoi(argv[4]) : 1;
	Loaded l;
	if(!loadScene(argv[2], l, threads)) return 1;
	session__.setStatusRenderStarted();
	auto t0 = std::chrono::steady_clock::now();
	bool ok = l.scene->render();
	auto t1 = std::chrono::steady_clock::now();
	if(!ok) { fprintf(stderr, "yaf_ref: render failed\n"); return 1; }
	ImageFilm *film = l.scene->getImageFilm();
	const int w = film->w_, h = film->h_;
	Rgba2DImageWeighed_t &img = *film->image_passes_[0];
	std::vector<float> buf((size_t)w * h * 5);
	for(int y = 0; y < h; ++y) for(int x = 0; x < w; ++x)
	{
		const Pixel &p = img(x, y);
		float *o = &buf[((size_t)y * w + x) * 5];
		o[0] = p.col_.r_; o[1] = p.col_.g_; o[2] = p.col_.b_; o[3] = p.col_.a_; o[4] = p.weight_;
	}
	FILE *f = fopen(argv[3], "wb");
	if(!f) { perror("open out"); return 1; }
	fwrite("YGF1", 1, 4, f);
	int32_t hdr[2] = {w, h};
	fwrite(hdr, sizeof(int32_t), 2, f);
	fwrite(buf.data(), sizeof(float), buf.size(), f);
	fclose(f);
	double wall = std::chrono::duration<double>(t1 - t0).count();
	// "rendert" is the reference's own timer around the passes (integrator_tiled.cc:168-169,253-255)
	printf("{\"width\": %d, \"height\": %d, \"threads\": %d, \"rendert_s\": %.6f, \"wall_s\": %.6f, \"ntris\": %d}\n",
	       w, h, l.scene->getNumThreads(), g_timer__.getTime("rendert"), wall, l.scene->tree_ ? (int)l.scene->tree_->total_prims_ : 0);
	return 0;
}

int cmdBench(int argc, char **argv)
{
	if(argc < 6) { fprintf(stderr, "usage: yaf_ref bench scene.xml threads warmup steps [spp]\n"); return 2; }
	const int threads = atoi(argv[3]), warmup = atoi(argv[4]), steps = atoi(argv[5]);
	const int spp = argc > 6 ? atoi(argv[6]) : 0;
	Loaded l;
	logger__.setConsoleMasterVerbosity("warning");
	logger__.setLogMasterVerbosity("mute");
	l.env = new RenderEnvironment();
	l.scene = new Scene(l.env);
	l.env->setScene(l.scene);
	auto t0 = std::chrono::steady_clock::now();
	if(!parseXmlFile__(argv[2], l.scene, l.env, l.render, "LinearRGB", 1.f)) return 1;
	auto t1 = std::chrono::steady_clock::now();
	l.render["threads"] = threads;
	l.render["logging_saveLog"] = false;
	l.render["logging_saveHTML"] = false;
	if(spp > 0) { l.render["AA_minsamples"] = spp; l.render["AA_passes"] = 1; }
	if(!l.env->setupScene(*l.scene, l.render, l.out)) return 1;
	session__.setInteractive(false);
	const std::map<std::string, Camera *> *cams = l.env->getCameraTable();
	if(cams->empty()) { fprintf(stderr, "no camera\n"); return 1; }
	l.scene->setCamera(cams->begin()->second);
	auto t2 = std::chrono::steady_clock::now();
	if(!l.scene->update()) { fprintf(stderr, "scene update failed\n"); return 1; }   // kd-tree build (kdtree_triangle.cc:76-157) + Light::init
	auto t3 = std::chrono::steady_clock::now();
	std::vector<double> rendert;
	session__.setStatusRenderStarted();
	for(int i = 0; i < warmup + steps; ++i)
	{
		if(!l.scene->render()) { fprintf(stderr, "yaf_ref: render failed\n"); return 1; }
		if(i >= warmup) rendert.push_back(g_timer__.getTime("rendert"));
	}
	ImageFilm *film = l.scene->getImageFilm();
	printf("{\"width\": %d, \"height\": %d, \"threads\": %d, \"spp\": %d, \"ntris\": %d, \"xml_parse_s\": %.3f, \"scene_update_s\": %.3f, \"rendert_s\": [",
	       film->w_, film->h_, l.scene->getNumThreads(), spp, l.scene->tree_ ? (int)l.scene->tree_->total_prims_ : 0,
	       std::chrono::duration<double>(t1 - t0).count(), std::chrono::duration<double>(t3 - t2).count());
	for(size_t i = 0; i < rendert.size(); ++i) printf("%s%.6f", i ? ", " : "", rendert[i]);
	printf("]}\n");
	return 0;
}

struct MeshRange { const Triangle *base; size_t count; size_t offset; };

int cmdHits(int argc, char **argv)
{
	if(argc < 5) { fprintf(stderr, "usage: yaf_ref hits scene.xml rays.bin hits.bin [mode] [shadow_depth]\n"); return 2; }
	int mode = argc > 5 ? atoi(argv[5]) : 0;
	int shadow_depth = argc > 6 ? atoi(argv[6]) : 2;
	Loaded l;
	if(!loadScene(argv[2], l, 1)) return 1;
	const std::map<std::string, Camera *> *cams = l.env->getCameraTable();
	if(cams->empty()) { fprintf(stderr, "no camera\n"); return 1; }
	l.scene->setCamera(cams->begin()->second);
	if(!l.scene->update()) { fprintf(stderr, "scene update failed\n"); return 1; }

	// global triangle order: scene.cc:807-817
	std::vector<MeshRange> ranges;
	size_t off = 0;
	for(auto &kv : l.scene->meshes_)
	{
		ObjData &dat = kv.second;
		if(!dat.obj_->isVisible() || dat.obj_->isBaseObject() || dat.type_ != 0 /*TRIM*/) continue;
		ranges.push_back({dat.obj_->triangles_.data(), dat.obj_->triangles_.size(), off});
		off += dat.obj_->triangles_.size();
	}
	auto globalId = [&](const Triangle *t) -> int32_t
	{
		for(const MeshRange &r : ranges) if(t >= r.base && t < r.base + r.count) return (int32_t)(r.offset + (t - r.base));
		return -2;
	};

	std::ifstream in(argv[3], std::ios::binary);
	std::vector<char> raw((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
	const size_t n = raw.size() / 32;
	const float *rays = (const float *)raw.data();
	struct Hit { float t, u, v; int32_t tri; };
	std::vector<Hit> hits(n);
	Random prng(1);
	RenderState state(&prng);
	state.time_ = 0.f;
	auto t0 = std::chrono::steady_clock::now();
	for(size_t i = 0; i < n; ++i)
	{
		const float *r = rays + i * 8;
		Ray ray(Point3(r[0], r[1], r[2]), Vec3(r[4], r[5], r[6]), r[3], r[7], 0.f);
		Hit &h = hits[i];
		h.t = 0.f; h.u = 0.f; h.v = 0.f; h.tri = -1;
		if(mode == 0)
		{
			SurfacePoint sp;
			if(l.scene->intersect(ray, sp))
			{
				h.t = ray.tmax_;
				h.u = sp.data_.b_1_; h.v = sp.data_.b_2_;
				h.tri = globalId((const Triangle *)sp.origin_);
			}
		}
		else if(mode == 1)
		{
			float oi = 0.f, mi = 0.f;
			// replicate isShadowed but recover the blocker triangle (scene.cc:962-981)
			Ray sray(ray);
			sray.from_ += sray.dir_ * sray.tmin_;
			float dis = (ray.tmax_ < 0) ? std::numeric_limits<float>::infinity() : sray.tmax_ - 2 * sray.tmin_;
			Triangle *hitt = nullptr;
			bool sh = l.scene->tree_->intersectS(sray, dis, &hitt, l.scene->shadow_bias_);
			bool sh2 = l.scene->isShadowed(state, ray, oi, mi);
			if(sh != sh2) { fprintf(stderr, "isShadowed mismatch\n"); return 1; }
			if(sh) { h.tri = globalId(hitt); h.t = 1.f; }
		}
		else
		{
			float oi = 0.f, mi = 0.f;
			Rgb filt(1.f);
			bool sh = l.scene->isShadowed(state, ray, shadow_depth, filt, oi, mi);
			h.tri = sh ? 0 : -1;
			h.t = filt.r_; h.u = filt.g_; h.v = filt.b_;
		}
	}
	auto t1 = std::chrono::steady_clock::now();
	FILE *f = fopen(argv[4], "wb");
	fwrite(hits.data(), sizeof(Hit), n, f);
	fclose(f);
	printf("{\"rays\": %zu, \"seconds\": %.6f, \"ntris\": %zu}\n", n, std::chrono::duration<double>(t1 - t0).count(), off);
	return 0;
}

// yaf_ref bgdump <scene.xml> <dirs.bin> <out.bin>: Background::eval for n directions (3 x float32 each) -> n x rgb float32
int cmdBgDump(int argc, char **argv)
{
	if(argc < 5) { fprintf(stderr, "usage: yaf_ref bgdump scene.xml dirs.bin out.bin\n"); return 2; }
	Loaded l;
	if(!loadScene(argv[2], l, 1)) return 1;
	const Background *bg = l.scene->getBackground();
	if(!bg) { fprintf(stderr, "no background\n"); return 1; }
	std::ifstream in(argv[3], std::ios::binary);
	std::vector<char> raw((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
	const size_t n = raw.size() / 12;
	const float *d = (const float *)raw.data();
	std::vector<float> out(n * 3);
	for(size_t i = 0; i < n; ++i)
	{
		Ray ray(Point3(0.f, 0.f, 0.f), Vec3(d[3 * i], d[3 * i + 1], d[3 * i + 2]), 0.f, -1.f, 0.f);
		const Rgb c = bg->eval(ray, false);
		out[3 * i] = c.r_; out[3 * i + 1] = c.g_; out[3 * i + 2] = c.b_;
	}
	FILE *f = fopen(argv[4], "wb");
	fwrite(out.data(), sizeof(float), out.size(), f);
	fclose(f);
	printf("{\"dirs\": %zu}\n", n);
	return 0;
}

} // namespace

#include "kat.h"

int main(int argc, char **argv)
{
	if(argc < 2) { fprintf(stderr, "usage: yaf_ref render|hits|kat ...\n"); return 2; }
	std::string cmd = argv[1];
	if(cmd == "render") return cmdRender(argc, argv);
	if(cmd == "hits") return cmdHits(argc, argv);
	if(cmd == "bench") return cmdBench(argc, argv);
	if(cmd == "kat") return cmdKat(argc, argv);
	if(cmd == "bgdump") return cmdBgDump(argc, argv);
	fprintf(stderr, "unknown command %s\n", argv[1]);
	return 2;
}
