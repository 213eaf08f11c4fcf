// TEST INFRASTRUCTURE — included by yaf_ref.cc.  `yaf_ref kat out.bin` dumps known-answer vectors produced by
// calling the reference's own helper functions on deterministic input grids.  tests/ regenerates the same grids
// (tests/kat_grids.py) and checks both the C restatement (oracle/yaf_oracle.c) and the CUDA device functions
// against these outputs.  Record format: u32 name_len, name, u32 dtype (0=f32 1=f64 2=i32 3=u32), u64 count, data.

BEGIN_YAFARAY
// non-static filter kernels of /root/reference/src/common/imagefilm.cc:60-122
float box__(float dx, float dy);
float mitchell__(float dx, float dy);
float gauss__(float dx, float dy);
float lanczos2__(float dx, float dy);
END_YAFARAY

struct KatWriter
{
	FILE *f;
	void put(const char *name, uint32_t dtype, const void *data, uint64_t count, size_t elem)
	{
		uint32_t nl = (uint32_t)strlen(name);
		fwrite(&nl, 4, 1, f); fwrite(name, 1, nl, f);
		fwrite(&dtype, 4, 1, f); fwrite(&count, 8, 1, f);
		fwrite(data, elem, count, f);
	}
	void f32(const char *n, const std::vector<float> &v) { put(n, 0, v.data(), v.size(), 4); }
	void f64(const char *n, const std::vector<double> &v) { put(n, 1, v.data(), v.size(), 8); }
	void u32(const char *n, const std::vector<uint32_t> &v) { put(n, 3, v.data(), v.size(), 4); }
};

// deterministic input generator shared with tests/kat_grids.py: 32-bit LCG -> float in [0,1)
static inline uint32_t katLcg(uint32_t &s) { s = s * 1664525u + 1013904223u; return s; }
static inline float katUnit(uint32_t &s) { return (float)(katLcg(s) >> 8) * (1.0f / 16777216.0f); }

static int cmdKat(int argc, char **argv)
{
	if(argc < 3) { fprintf(stderr, "usage: yaf_ref kat out.bin\n"); return 2; }
	KatWriter w{fopen(argv[2], "wb")};
	if(!w.f) { perror("open"); return 1; }
	const int n = 4096;

	{	// integer hash + radical inverses (util_mcqmc.h:93-162)
		std::vector<uint32_t> h(n);
		std::vector<float> vdc(n), sob(n), lp(n), vdc0(n), sob0(n);
		for(int i = 0; i < n; ++i)
		{
			uint32_t x = (uint32_t)i * 2654435761u + (uint32_t)i;
			h[i] = fnv32ABuf__(x);
			vdc[i] = riVdC__(x, h[i]); sob[i] = riS__(x, h[i]); lp[i] = riLp__(x + h[i]);
			vdc0[i] = riVdC__((uint32_t)i); sob0[i] = riS__((uint32_t)i);
		}
		w.u32("fnv32", h); w.f32("riVdC", vdc); w.f32("riS", sob); w.f32("riLp", lp); w.f32("riVdC0", vdc0); w.f32("riS0", sob0);
	}
	{	// scrambled Halton (scr_halton.h:52-76): dims 1..49, n = k*7919+dim for k<64 plus huge offsets
		std::vector<double> v;
		for(int dim = 1; dim < 50; ++dim) for(uint32_t k = 0; k < 64; ++k)
		{
			v.push_back(scrHalton__(dim, k * 7919u + (uint32_t)dim));
			v.push_back(scrHalton__(dim, 0xfff00000u + k * 104729u));
		}
		w.f64("scrHalton", v);
	}
	{	// incremental Halton (util_mcqmc.h:28-85): bases 2,3,5,7; start values incl. the unsigned wrap of (0u-1)
		std::vector<float> v;
		const uint32_t starts[] = {0u, 1u, 7u, 4566u, 123456789u, 0xffffffffu, 0x9e3779b9u};
		const int bases[] = {2, 3, 5, 7};
		for(int b : bases) for(uint32_t s : starts)
		{
			Halton hal(b);
			hal.setStart(s);
			for(int i = 0; i < 16; ++i) v.push_back(hal.getNext());
		}
		w.f32("halton", v);
	}
	{	// MWC generator (util_mcqmc.h:173-190)
		std::vector<double> v;
		const uint32_t seeds[] = {0u, 1u, 123u, 1804289383u + 123u};
		for(uint32_t s : seeds) { Random r(s); for(int i = 0; i < 64; ++i) v.push_back(r()); }
		w.f64("random", v);
	}
	{	// CIE 1931 colour matching functions as the reference tabulates them (spectrum.cc): wl2Xyz__ at the integer wavelengths 360..829 nm
		// returns the table rows themselves (fr = 0); row 470 (830 nm) is recovered from the midpoint 829.5 nm.  Also a few wl2Rgb__ / getIor__ /
		// cauchyCoefficients__ known answers for the dispersion code.
		std::vector<float> t(471 * 3, 0.f);
		for(int i = 0; i < 470; ++i) { const Rgb c = wl2Xyz__(360.f + (float)i); t[3 * i] = c.r_; t[3 * i + 1] = c.g_; t[3 * i + 2] = c.b_; }
		{ const Rgb c = wl2Xyz__(829.5f); for(int k = 0; k < 3; ++k) t[3 * 470 + k] = 2.f * (k == 0 ? c.r_ : k == 1 ? c.g_ : c.b_) - t[3 * 469 + k]; }
		w.f32("cie_xyz", t);
		std::vector<float> v;
		for(int i = 0; i < 64; ++i)
		{
			const float wl = (float)i / 64.f + 0.0078125f;
			Rgb c; wl2Rgb__(wl, c);
			float ca, cb; cauchyCoefficients__(1.3f + 0.01f * (float)i, 5.f + (float)i, ca, cb);
			v.push_back(c.r_); v.push_back(c.g_); v.push_back(c.b_); v.push_back(ca); v.push_back(cb); v.push_back(getIor__(wl, ca, cb));
		}
		w.f32("dispersion_kat", v);
	}
	{	// fast math (util_math_optimizations.h:116-270)
		std::vector<float> fsin(n), fcos(n), fexp(n), flog(n), fpow(n), fexp2(n), flog2(n), facos(n), fasin(n), fsqrt(n);
		uint32_t s = 12345u;
		for(int i = 0; i < n; ++i)
		{
			float a = (katUnit(s) - 0.5f) * 40.f;       // trig argument in [-20,20]
			float b = katUnit(s) * 16.f + 1e-6f;          // positive
			float c = (katUnit(s) - 0.5f) * 2.2f;         // around [-1.1,1.1]
			float e = katUnit(s) * 200.f;                 // exponent
			fsin[i] = fSin__(a); fcos[i] = fCos__(a);
			fexp[i] = fExp__(-b); flog[i] = fLog__(b); fexp2[i] = fExp2__(a); flog2[i] = fLog2__(b);
			fpow[i] = fPow__(katUnit(s), e);
			facos[i] = fAcos__(c); fasin[i] = fAsin__(c); fsqrt[i] = fSqrt__(b);
		}
		w.f32("fSin", fsin); w.f32("fCos", fcos); w.f32("fExp", fexp); w.f32("fLog", flog); w.f32("fExp2", fexp2);
		w.f32("fLog2", flog2); w.f32("fPow", fpow); w.f32("fAcos", facos); w.f32("fAsin", fasin); w.f32("fSqrt", fsqrt);
	}
	{	// sampling helpers (util_sample.h:45-86,183-187) and createCs__ (vector.h:319-337)
		std::vector<float> v, cs, cone, am;
		uint32_t s = 777u;
		for(int i = 0; i < 1024; ++i)
		{
			Vec3 nn(katUnit(s) - 0.5f, katUnit(s) - 0.5f, katUnit(s) - 0.5f);
			if(i % 97 == 0) { nn.x_ = 0.f; nn.y_ = 0.f; }
			nn.normalize();
			Vec3 u, vv;
			createCs__(nn, u, vv);
			float s1 = katUnit(s), s2 = katUnit(s);
			if(i % 101 == 0) s1 = 1.0f;
			Vec3 d = sampleCosHemisphere__(nn, u, vv, s1, s2);
			Vec3 c = sampleCone__(nn, u, vv, 0.95f + 0.05f * katUnit(s), s1, s2);
			cs.insert(cs.end(), {nn.x_, nn.y_, nn.z_, u.x_, u.y_, u.z_, vv.x_, vv.y_, vv.z_});
			v.insert(v.end(), {d.x_, d.y_, d.z_});
			cone.insert(cone.end(), {c.x_, c.y_, c.z_});
			am.push_back(addMod1__(s1, s2));
		}
		w.f32("createCs", cs); w.f32("sampleCosHemisphere", v); w.f32("sampleCone", cone); w.f32("addMod1", am);
	}
	{	// refract / fresnel (src/common/vector.cc:86-141)
		std::vector<float> v;
		uint32_t s = 4242u;
		for(int i = 0; i < 1024; ++i)
		{
			Vec3 nn(katUnit(s) - 0.5f, katUnit(s) - 0.5f, katUnit(s) - 0.5f); nn.normalize();
			Vec3 wi(katUnit(s) - 0.5f, katUnit(s) - 0.5f, katUnit(s) - 0.5f); wi.normalize();
			float ior = 1.f + katUnit(s);
			Vec3 wo(0.f);
			bool ok = refract__(nn, wi, wo, ior);
			float kr = 0.f, kt = 0.f;
			fresnel__(wi, nn, ior, kr, kt);
			v.insert(v.end(), {ok ? 1.f : 0.f, ok ? wo.x_ : 0.f, ok ? wo.y_ : 0.f, ok ? wo.z_ : 0.f, kr, kt});
		}
		w.f32("refract_fresnel", v);
	}
	{	// filter tables exactly as ImageFilm::ImageFilm fills them (imagefilm.cc:158-176)
		float (*fn[4])(float, float) = {box__, mitchell__, gauss__, lanczos2__};
		std::vector<float> v;
		const float scale = 1.f / 16.f;
		for(int k = 0; k < 4; ++k) for(int y = 0; y < 16; ++y) for(int x = 0; x < 16; ++x) v.push_back(fn[k]((x + .5f) * scale, (y + .5f) * scale));
		w.f32("filter_tables", v);
	}
	fclose(w.f);
	return 0;
}
