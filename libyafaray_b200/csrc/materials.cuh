// K5 + K7 device code: surface-point construction and the per-material BSDF functions.
//
// Restates, operation for operation (float32, no FMA contraction), the constant-parameter (no shader node) paths of
//   Triangle::getSurface            /root/reference/src/common/triangle.cc:30-133
//   ShinyDiffuseMaterial            /root/reference/src/material/material_shiny_diffuse.cc:98-597
//   GlassMaterial / Mirror / Null   /root/reference/src/material/material_glass.cc:51-309,466-500
//   LightMaterial                   /root/reference/src/material/material_simple.cc:38-61
//   GlossyMaterial                  /root/reference/src/material/material_glossy.cc:51-405
//   CoatedGlossyMaterial            /root/reference/src/material/material_coated_glossy.cc:67-463
//   RoughGlassMaterial              /root/reference/src/material/material_rough_glass.cc:50-320
//   BlendMaterial                   /root/reference/src/material/material_blend.cc:57-462
//   microfacet helpers              /root/reference/include/material/material_utils_microfacet.h:35-210
// The per-hit "userdata" scratch of the reference (SdDat / MDatT) holds only values that are constants of the material
// when no shader nodes exist, so it is folded into MatDev by host_prep.h.
#pragma once
#include "scene_dev.cuh"

namespace yg {

enum : unsigned
{
	BsdfNone = 0x0000, BsdfSpecular = 0x0001, BsdfGlossy = 0x0002, BsdfDiffuse = 0x0004, BsdfDispersive = 0x0008,
	BsdfReflect = 0x0010, BsdfTransmit = 0x0020, BsdfFilter = 0x0040, BsdfEmit = 0x0080, BsdfVolumetric = 0x0100,
	BsdfAllSpecular = BsdfSpecular | BsdfReflect | BsdfTransmit,
	BsdfAllGlossy = BsdfGlossy | BsdfReflect | BsdfTransmit,
	BsdfAll = BsdfSpecular | BsdfGlossy | BsdfDiffuse | BsdfDispersive | BsdfReflect | BsdfTransmit | BsdfFilter
}; // material.h:50-66

struct SurfPt
{
	V3 p, n, ng, nu, nv;
	int mat;    // material index
	int light;  // light index the hit object is tagged with (ObjectGeometric::light_), -1 = none
	int tri;    // global triangle id
	// RenderState::chromatic_ / wavelength_ (scene.h:97-99) of the task being shaded: only dispersive glass reads them
	int chromatic; float wavelength;
};

// getIor__ (spectrum.h:35-39): wl = 300.0 * w + 400.0 is a double expression narrowed to float
YG_D float cauchyIor(float w, float cauchy_a, float cauchy_b)
{
	const float wl = (float)(300.0 * (double)w + 400.0);
	return cauchy_a + cauchy_b / (wl * wl);
}
// IOR the glass materials use at this hit: the Cauchy value at the task's wavelength while the full spectrum is still being carried
// (disperse_ && state.chromatic_, material_glass.cc:83-96,148-155,267-274, material_rough_glass.cc:85-92,199-206), else ior_
#define YG_IOR(m, sp) (((m).disperse && (sp).chromatic) ? cauchyIor((sp).wavelength, (m).cauchy_a, (m).cauchy_b) : (m).ior)

struct BSample { float s1, s2, pdf; unsigned flags, sampled; };

// wl2Rgb__ (spectrum.h:41-46) = wl2RgbFromCie__ (spectrum.cc:276-289: linear interpolation in the CIE 1931 table) + xyzToRgb__ (:264-273)
YG_D Col wl2Rgb(const SceneDev &sc, float w)
{
	const float wl = (float)(300.0 * (double)w + 400.0);
	float fr = wl - 360.f;
	const int p_1 = (int)fr;
	if(p_1 < 0 || p_1 + 1 > 470 || !sc.cie_xyz) return mkc(0.f);
	fr -= floorf(fr);
	const float fr_2 = 1.f - fr;
	const float *a = sc.cie_xyz + 3 * p_1, *b = a + 3;
	const float x = fr_2 * __ldg(a) + fr * __ldg(b), y = fr_2 * __ldg(a + 1) + fr * __ldg(b + 1), z = fr_2 * __ldg(a + 2) + fr * __ldg(b + 2);
	Col c = mkc((2.28783848734076f * x - 0.833367677835217f * y) - 0.454470795871421f * z,
	            (-0.511651380743862f * x + 1.42275837632178f * y) + 0.0888930017552939f * z,
	            (0.00572040983140966f * x - 0.0159068485104036f * y) + 1.0101864083734f * z);
	const float wt = fminf(c.r, fminf(c.g, c.b));
	if(wt < 0.f) c = mkc(c.r - wt, c.g - wt, c.b - wt);
	return c * 2.214032659670777114f;
}

YG_HD V3 ldv(const float *p) { return mk(p[0], p[1], p[2]); }
YG_HD Col ldc(const float *p) { return mkc(p[0], p[1], p[2]); }

// Triangle::getSurface triangle.cc:30-133 (geometry part; orco/uv/dPdU feed only shader nodes, which are out of scope)
YG_D void getSurface(const SceneDev &sc, int tri, float b1, float b2, V3 hit, SurfPt &sp)
{
	const TriShade *ts = sc.tris + tri;
	const int4 q0 = __ldg((const int4 *)ts), q1 = __ldg((const int4 *)ts + 1), q2 = __ldg((const int4 *)ts + 2);
	const int2 q3 = __ldg((const int2 *)((const int4 *)ts + 3));
	// layout: v[3] n[3] uv[3] mat obj ng[3]
	const int n0 = q0.w, n1 = q1.x, n2 = q1.y;
	const int mat = q2.y, obj = q2.z;
	sp.ng = mk(__int_as_float(q2.w), __int_as_float(q3.x), __int_as_float(q3.y));
	const ObjDev ob = sc.objects[obj];
	const float u = (1 - b1) - b2, v = b1, w = b2; // data.b_0_, b_1_, b_2_ (triangle.h:252-254, triangle.cc:34)
	if(ob.smooth)
	{
		const V3 va = n0 >= 0 ? ldv(sc.normals + 3 * n0) : sp.ng;
		const V3 vb = n1 >= 0 ? ldv(sc.normals + 3 * n1) : sp.ng;
		const V3 vc = n2 >= 0 ? ldv(sc.normals + 3 * n2) : sp.ng;
		sp.n = normalize((u * va + v * vb) + w * vc);
	}
	else sp.n = sp.ng;
	sp.p = hit;
	createCs(sp.n, sp.nu, sp.nv);
	sp.mat = mat; sp.light = ob.light; sp.tri = tri; sp.chromatic = 1; sp.wavelength = 0.f;
}

// ---- helpers --------------------------------------------------------------------------------------------------------
#define YG_SDL(MT) ((MT) == YG_MAT_SD_LAMBERT) /* compile-time: the kernel of the diffuse-only shinydiffuse specialisation */

// ShinyDiffuseMaterial::getFresnel material_shiny_diffuse.cc:120-148
template<int MT>
YG_D float sdFresnel(const MatDev &m, V3 wo, V3 n)
{
	if(YG_SDL(MT) || !m.has_fresnel) return 1.f;
	const V3 N = (dot(wo, n) < 0.f) ? -n : n;
	const float c = dot(wo, N);
	float g = m.ior_squared + c * c - 1.f;
	if(g < 0.f) g = 0.f; else g = fSqrt(g);
	const float aux = c * (g + c);
	return ((0.5f * (g - c) * (g - c)) / ((g + c) * (g + c))) * (1.f + ((aux - 1) * (aux - 1)) / ((aux + 1) * (aux + 1)));
}
// accumulate__ :153-162
YG_D void sdAccumulate(const float *component, float *accum, float kr)
{
	accum[0] = component[0] * kr;
	float acc = 1.f - accum[0];
	accum[1] = component[1] * acc;
	acc *= 1.f - component[1];
	accum[2] = component[2] * acc;
	acc *= 1.f - component[2];
	accum[3] = component[3] * acc;
}
template<int MT>
YG_D void sdComponents(const MatDev &m, float *c)
{	// initBsdf: memset 0 then getComponents without nodes (:98-118,164-183)
	c[0] = (!YG_SDL(MT) && m.is_mirror) ? m.mirror_strength : 0.f;
	c[1] = (!YG_SDL(MT) && m.is_transparent) ? m.transparency_strength : 0.f;
	c[2] = (!YG_SDL(MT) && m.is_translucent) ? m.translucency_strength : 0.f;
	c[3] = (YG_SDL(MT) || m.is_diffuse) ? m.diffuse_strength : 0.f;
}
// component table of config() (:46-92); constants for the Lambert specialisation
template<int MT> YG_D int sdNBsdf(const MatDev &m) { return YG_SDL(MT) ? 1 : m.n_bsdf; }
template<int MT> YG_D unsigned sdCFlags(const MatDev &m, int i) { return YG_SDL(MT) ? (unsigned)(BsdfDiffuse | BsdfReflect) : m.c_flags[i]; }
template<int MT> YG_D int sdCIndex(const MatDev &m, int i) { return YG_SDL(MT) ? 3 : m.c_index[i]; }
// orenNayar (same code in shiny diffuse :204-241, glossy :74-111, coated glossy :92-128)
YG_D float orenNayar(float oren_a, float oren_b, V3 wi, V3 wo, V3 n)
{
	const float cos_ti = fmaxf(-1.f, fminf(1.f, dot(n, wi)));
	const float cos_to = fmaxf(-1.f, fminf(1.f, dot(n, wo)));
	float maxcos_f = 0.f;
	if(cos_ti < 0.9999f && cos_to < 0.9999f)
	{
		const V3 v1 = normalize(wi - n * cos_ti);
		const V3 v2 = normalize(wo - n * cos_to);
		maxcos_f = fmaxf(0.f, dot(v1, v2));
	}
	float sin_alpha, tan_beta;
	if(cos_to >= cos_ti)
	{
		sin_alpha = fSqrt(1.f - cos_ti * cos_ti);
		tan_beta = fSqrt(1.f - cos_to * cos_to) / ((cos_to == 0.f) ? 1e-8f : cos_to);
	}
	else
	{
		sin_alpha = fSqrt(1.f - cos_to * cos_to);
		tan_beta = fSqrt(1.f - cos_ti * cos_ti) / ((cos_ti == 0.f) ? 1e-8f : cos_ti);
	}
	return fminf(1.f, fmaxf(0.f, oren_a + oren_b * maxcos_f * sin_alpha * tan_beta));
}

// ShinyDiffuseMaterial::orenNayar with use_texture_sigma (material_shiny_diffuse.cc:229-236): coefficients and the final sum in double
YG_D float orenNayarTex(float sigma, V3 wi, V3 wo, V3 n)
{
	const float cos_ti = fmaxf(-1.f, fminf(1.f, dot(n, wi)));
	const float cos_to = fmaxf(-1.f, fminf(1.f, dot(n, wo)));
	float maxcos_f = 0.f;
	if(cos_ti < 0.9999f && cos_to < 0.9999f)
	{
		const V3 v1 = normalize(wi - n * cos_ti);
		const V3 v2 = normalize(wo - n * cos_to);
		maxcos_f = fmaxf(0.f, dot(v1, v2));
	}
	float sin_alpha, tan_beta;
	if(cos_to >= cos_ti)
	{
		sin_alpha = fSqrt(1.f - cos_ti * cos_ti);
		tan_beta = fSqrt(1.f - cos_to * cos_to) / ((cos_to == 0.f) ? 1e-8f : cos_to);
	}
	else
	{
		sin_alpha = fSqrt(1.f - cos_to * cos_to);
		tan_beta = fSqrt(1.f - cos_ti * cos_ti) / ((cos_ti == 0.f) ? 1e-8f : cos_ti);
	}
	const double s2 = (double)sigma * (double)sigma;
	const double a = 1.0 - 0.5 * (s2 / (s2 + 0.33)), b = 0.45 * s2 / (s2 + 0.09);
	return fminf(1.f, fmaxf(0.f, (float)(a + b * (double)maxcos_f * (double)sin_alpha * (double)tan_beta)));
}

// microfacet helpers, material_utils_microfacet.h
YG_D float pdfDivisor(float c) { return (float)((double)8.f * YG_M_PI * (double)(c * 0.99f + 0.04f)); }
YG_D float asDivisor(float c1, float ci, float co) { return (float)((double)8.f * YG_M_PI * (double)((c1 * fmaxf(ci, co)) * 0.99f + 0.04f)); }
// the reference divides a float by the double-valued macro: x / DIVISOR is a double division rounded to float
YG_D float divPdfDivisor(float x, float c) { return (float)((double)x / ((double)8.f * YG_M_PI * (double)(c * 0.99f + 0.04f))); }
YG_D float divAsDivisor(float x, float c1, float ci, float co) { return (float)((double)x / ((double)8.f * YG_M_PI * (double)((c1 * fmaxf(ci, co)) * 0.99f + 0.04f))); }
YG_D float blinnD(float cos_h, float e) { return (e + 1.f) * fPow(cos_h, e); }
YG_D float blinnPdf(float costheta, float cos_w_h, float e) { return divPdfDivisor(blinnD(costheta, e), cos_w_h); }
YG_D V3 blinnSample(float s1, float s2, float exponent)
{
	const float cos_theta = fPow(1.f - s2, 1.f / (exponent + 1.f));
	const float sin_theta = fSqrt(1.f - cos_theta * cos_theta);
	const float phi = mulD(s1, YG_M_2PI);
	return mk(sin_theta * fCos(phi), sin_theta * fSin(phi), cos_theta);
}
YG_D float asAnisoD(V3 h, float e_u, float e_v)
{
	if(h.z <= 0.f) return 0.f;
	const float exponent = (e_u * h.x * h.x + e_v * h.y * h.y) / (1.00001f - h.z * h.z);
	return fSqrt((e_u + 1.f) * (e_v + 1.f)) * fPow(fmaxf(0.f, h.z), exponent);
}
YG_D float asAnisoPdf(V3 h, float cos_w_h, float e_u, float e_v) { return divPdfDivisor(asAnisoD(h, e_u, e_v), cos_w_h); }
YG_D V3 sampleQuadrantAniso(float s1, float s2, float e_u, float e_v)
{
	const float phi = atanf(fSqrt((e_u + 1.f) / (e_v + 1.f)) * tanf(mulD(s1, 1.57079632679489661923)));
	const float cos_phi = fCos(phi), sin_phi = fSin(phi);
	const float cos_phi_2 = cos_phi * cos_phi, sin_phi_2 = 1.f - cos_phi_2;
	const float cos_theta = fPow(1.f - s2, 1.f / (e_u * cos_phi_2 + e_v * sin_phi_2 + 1.f));
	const float sin_theta = fSqrt(1.f - cos_theta * cos_theta);
	return mk(sin_theta * cos_phi, sin_theta * sin_phi, cos_theta);
}
YG_D V3 asAnisoSample(float s1, float s2, float e_u, float e_v)
{
	V3 h;
	if(s1 < 0.25f) h = sampleQuadrantAniso(4.f * s1, s2, e_u, e_v);
	else if(s1 < 0.5f) { h = sampleQuadrantAniso(1.f - 4.f * (0.5f - s1), s2, e_u, e_v); h.x = -h.x; }
	else if(s1 < 0.75f) { h = sampleQuadrantAniso(4.f * (s1 - 0.5f), s2, e_u, e_v); h.x = -h.x; h.y = -h.y; }
	else { h = sampleQuadrantAniso(1.f - 4.f * (1.f - s1), s2, e_u, e_v); h.y = -h.y; }
	return h;
}
YG_D float schlickFresnel(float costheta, float r)
{
	const float c1 = (1.f - costheta);
	const float c2 = c1 * c1;
	return r + ((1.f - r) * c1 * c2 * c2);
}
YG_D Col diffuseReflect(float wi_n, float wo_n, float glossy, float diffuse, Col diff_base)
{
	float f_wi = (1.f - (0.5f * wi_n));
	float temp = f_wi * f_wi;
	f_wi = temp * temp * f_wi;
	float f_wo = (1.f - (0.5f * wo_n));
	temp = f_wo * f_wo;
	f_wo = temp * temp * f_wo;
	// DIFFUSE_RATIO is a double literal: the scalar chain runs in double and narrows at Rgb operator*(float, Rgb)
	const float k = (float)(0.387507688 * (double)diffuse * (double)(1.f - glossy) * (double)(1.f - f_wi) * (double)(1.f - f_wo));
	return k * diff_base;
}
YG_D V3 ggxSample(float alpha_2, float s1, float s2)
{
	const float tan_theta_2 = alpha_2 * (s1 / (1.00001f - s1));
	const float cos_theta = 1.f / fSqrt(1.f + tan_theta_2);
	const float sin_theta = fSqrt(1.00001f - (cos_theta * cos_theta));
	const float phi = (float)(YG_M_2PI * (double)s2);
	return mk(sin_theta * fCos(phi), sin_theta * fSin(phi), cos_theta);
}
YG_D float ggxD(float alpha_2, float cos_theta_2, float tan_theta_2)
{
	const float cos_theta_4 = cos_theta_2 * cos_theta_2;
	const float a_tan = alpha_2 + tan_theta_2;
	const float div = (float)(YG_M_PI * (double)cos_theta_4 * (double)a_tan * (double)a_tan);
	return alpha_2 / div;
}
YG_D float ggxG(float alpha_2, float wo_n, float wi_n)
{
	const float wo_n_2 = wo_n * wo_n, wi_n_2 = wi_n * wi_n;
	const float t1 = fSqrt(1.f + alpha_2 * ((1.f - wo_n_2) / wo_n_2));
	const float t2 = fSqrt(1.f + alpha_2 * ((1.f - wi_n_2) / wi_n_2));
	return (2.f / (1.f + t1)) * (2.f / (1.f + t2));
}
YG_D float microfacetFresnel(float wo_h, float ior)
{
	const float c = fabsf(wo_h);
	float g = ior * ior - 1 + c * c;
	if(g > 0)
	{
		g = fSqrt(g);
		const float a = (g - c) / (g + c);
		const float b = (c * (g + c) - 1) / (c * (g - c) + 1);
		return 0.5f * a * a * (1 + b * b);
	}
	return 1.0f;
}
YG_D bool refractMicrofacet(float eta, V3 wo, V3 &wi, V3 h, float wo_h, float &kr, float &kt)
{
	wi = mk(0.f);
	const float c = dot(-wo, h);
	const float sign = (c > 0.f) ? 1.f : -1.f;
	const float t1 = 1 - (eta * eta * (1 - c * c));
	if(t1 < 0.f) return false;
	wi = eta * wo + (eta * c - sign * fSqrt(t1)) * h;
	wi = -wi;
	kr = 0.f; kt = 0.f;
	kr = microfacetFresnel(wo_h, 1.f / eta);
	if(kr == 1.f) return false;
	kt = 1 - kr;
	return true;
}
YG_D V3 reflectMicrofacet(V3 wo, V3 h)
{
	V3 wi = wo + (2.f * dot(h, -wo)) * h;
	return -wi;
}

// ---- leaf materials -------------------------------------------------------------------------------------------------

// glass normal fix-up shared by GlassMaterial::sample/getSpecular and CoatedGlossy::getSpecular
YG_D V3 glassNormal(const SurfPt &sp, V3 wo, bool outside)
{
	const float cos_wo_n = dot(sp.n, wo);
	const float k = (float)(1.00001 * (double)cos_wo_n);
	if(outside) return (cos_wo_n >= 0) ? sp.n : normalize(sp.n - k * wo);
	return (cos_wo_n <= 0) ? sp.n : normalize(sp.n - k * wo);
}

// MT >= 0: the material type is a compile-time constant (per-material shading kernels, integrator.cu); MT < 0: read it from the record
// YG_MAT_SD_LAMBERT (internal, host_prep.h): a shinydiffuse material whose only component is diffuse reflection and which has no
// Fresnel term.  It runs the SAME shinydiffuse code, but the component flags are compile-time constants in its kernel, so the
// component loops, the Fresnel/accumulate arithmetic on exact zeros and ones, and the dynamically indexed scratch arrays fold away.
template<int MT> YG_D int mtypeOf(const MatDev &m)
{
	const int t = MT >= 0 ? MT : m.type;
	return t == YG_MAT_SD_LAMBERT ? YAFGPU_MAT_SHINYDIFFUSE : t;
}


template<int MT>
YG_D Col leafEval(const MatDev &m, const SurfPt &sp, V3 wo, V3 wl, unsigned bsdfs)
{
	switch(mtypeOf<MT>(m))
	{
		case YAFGPU_MAT_SHINYDIFFUSE:
		{	// material_shiny_diffuse.cc:244-293
			const float cos_ng_wo = dot(sp.ng, wo), cos_ng_wl = dot(sp.ng, wl);
			const V3 n = faceForward(sp.ng, sp.n, wo);
			if(!(bsdfs & m.bsdf_flags & BsdfDiffuse)) return mkc(0.f);
			float comp[4]; sdComponents<MT>(m, comp);
			const float kr = sdFresnel<MT>(m, wo, n);
			const float m_t = (1.f - kr * comp[0]) * (1.f - comp[1]);
			const bool transmit = (cos_ng_wo * cos_ng_wl) < 0.f;
			if(!YG_SDL(MT) && transmit && m.is_translucent) return (comp[2] * m_t) * ldc(m.diffuse_col);
			if(dot(n, wl) < 0.0f && !m.flat) return mkc(0.f);
			float m_d = m_t * (1.f - comp[2]) * comp[3];
			if(m.use_oren_nayar) m_d *= (MT < 0 && m.oren_tex) ? orenNayarTex(m.oren_sigma, wo, wl, n) : orenNayar(m.oren_a, m.oren_b, wo, wl, n);
			if(MT < 0) m_d *= m.diffuse_refl_mul; // "diffuse_refl_shader" (:285); 1 when absent.  Only the generic instantiation can see nodes.
			return m_d * ldc(m.diffuse_col);
		}
		case YAFGPU_MAT_GLOSSY:
		{	// material_glossy.cc:113-173
			if(!(bsdfs & BsdfDiffuse) || (dot(sp.ng, wl) * dot(sp.ng, wo)) < 0.f) return mkc(0.f);
			Col col = mkc(0.f);
			const bool diffuse_flag = bsdfs & BsdfDiffuse;
			const V3 n = faceForward(sp.ng, sp.n, wo);
			const float wi_n = fabsf(dot(wl, n)), wo_n = fabsf(dot(wo, n));
			if((m.as_diffuse && diffuse_flag) || (!m.as_diffuse && (bsdfs & BsdfGlossy)))
			{
				const V3 h = normalize(wo + wl);
				const float cos_wi_h = fmaxf(0.f, dot(wl, h));
				float glossy;
				if(m.anisotropic)
				{
					const V3 hs = mk(dot(h, sp.nu), dot(h, sp.nv), dot(h, n));
					glossy = divAsDivisor(asAnisoD(hs, m.exp_u, m.exp_v) * schlickFresnel(cos_wi_h, m.gl_glossy), cos_wi_h, wo_n, wi_n);
				}
				else glossy = divAsDivisor(blinnD(dot(h, n), m.gl_exponent) * schlickFresnel(cos_wi_h, m.gl_glossy), cos_wi_h, wo_n, wi_n);
				col = glossy * ldc(m.gloss_col);
			}
			if(m.with_diffuse && diffuse_flag)
			{
				Col add_col = (m.gl_diffuse * (1.f - m.gl_glossy)) * ldc(m.diff_col);
				if(MT < 0) add_col = add_col * m.diffuse_refl_mul; // "diffuse_refl_shader"; 1 when absent
				if(m.gl_oren_nayar) add_col = add_col * ((MT < 0 && m.oren_tex) ? orenNayarTex(m.oren_sigma, wl, wo, n) : orenNayar(m.gl_oren_a, m.gl_oren_b, wl, wo, n));
				col = col + add_col;
			}
			return col;
		}
		case YAFGPU_MAT_COATED_GLOSSY:
		{	// material_coated_glossy.cc:130-185
			Col col = mkc(0.f);
			const bool diffuse_flag = bsdfs & BsdfDiffuse;
			if(!diffuse_flag || (dot(sp.ng, wl) * dot(sp.ng, wo)) < 0.f) return col;
			const V3 n = faceForward(sp.ng, sp.n, wo);
			float kr, kt;
			const float wi_n = fabsf(dot(wl, n)), wo_n = fabsf(dot(wo, n));
			fresnel(wo, n, m.ior, kr, kt);
			if((m.as_diffuse && diffuse_flag) || (!m.as_diffuse && (bsdfs & BsdfGlossy)))
			{
				const V3 h = normalize(wo + wl);
				const float cos_wi_h = dot(wl, h);
				float glossy;
				if(m.anisotropic)
				{
					const V3 hs = mk(dot(h, sp.nu), dot(h, sp.nv), dot(h, n));
					glossy = divAsDivisor(kt * asAnisoD(hs, m.exp_u, m.exp_v) * schlickFresnel(cos_wi_h, m.gl_glossy), cos_wi_h, wo_n, wi_n);
				}
				else glossy = divAsDivisor(kt * blinnD(dot(h, n), m.gl_exponent) * schlickFresnel(cos_wi_h, m.gl_glossy), cos_wi_h, wo_n, wi_n);
				col = glossy * ldc(m.gloss_col);
			}
			if(m.with_diffuse && diffuse_flag)
			{
				Col add_col = ((m.gl_diffuse * (1.f - m.gl_glossy)) * ldc(m.diff_col)) * kt;
				if(MT < 0) add_col = add_col * m.diffuse_refl_mul; // "diffuse_refl_shader"; 1 when absent
				if(m.gl_oren_nayar) add_col = add_col * ((MT < 0 && m.oren_tex) ? orenNayarTex(m.oren_sigma, wl, wo, n) : orenNayar(m.gl_oren_a, m.gl_oren_b, wl, wo, n));
				col = col + add_col;
			}
			return col;
		}
		default: return mkc(0.f); // glass, rough glass, mirror, null, light_mat
	}
}

template<int MT>
YG_D float leafPdf(const MatDev &m, const SurfPt &sp, V3 wo, V3 wi, unsigned bsdfs)
{
	switch(mtypeOf<MT>(m))
	{
		case YAFGPU_MAT_SHINYDIFFUSE:
		{	// material_shiny_diffuse.cc:410-460
			if(!(bsdfs & BsdfDiffuse)) return 0.f;
			float pdf = 0.f, accum_c[4], comp[4];
			const float cos_ng_wo = dot(sp.ng, wo);
			const V3 n = faceForward(sp.ng, sp.n, wo);
			sdComponents<MT>(m, comp);
			sdAccumulate(comp, accum_c, sdFresnel<MT>(m, wo, n));
			float sum = 0.f; int n_match = 0;
			for(int i = 0; i < sdNBsdf<MT>(m); ++i)
			{
				if(bsdfs & sdCFlags<MT>(m, i))
				{
					const float width = accum_c[sdCIndex<MT>(m, i)];
					sum += width;
					if(sdCFlags<MT>(m, i) == (BsdfDiffuse | BsdfTransmit))
					{
						if(cos_ng_wo * dot(sp.ng, wi) < 0) pdf += fabsf(dot(wi, n)) * width;
					}
					else if(sdCFlags<MT>(m, i) == (BsdfDiffuse | BsdfReflect)) pdf += fabsf(dot(wi, n)) * width;
					++n_match;
				}
			}
			if(!n_match || sum < 0.00001f) return 0.f;
			return pdf / sum;
		}
		case YAFGPU_MAT_GLOSSY:
		{	// material_glossy.cc:359-405
			if(dot(sp.ng, wo) * dot(sp.ng, wi) < 0.f) return 0.f;
			const V3 n = faceForward(sp.ng, sp.n, wo);
			float pdf = 0.f;
			const bool use_glossy = m.as_diffuse ? (bsdfs & BsdfDiffuse) : (bsdfs & BsdfGlossy);
			const bool use_diffuse = m.with_diffuse && (bsdfs & BsdfDiffuse);
			if(use_diffuse)
			{
				pdf = fabsf(dot(wi, n));
				if(use_glossy)
				{
					const V3 h = normalize(wi + wo);
					const float cos_wo_h = dot(wo, h), cos_n_h = dot(n, h);
					if(m.anisotropic) pdf = pdf * m.p_diffuse + asAnisoPdf(mk(dot(h, sp.nu), dot(h, sp.nv), cos_n_h), cos_wo_h, m.exp_u, m.exp_v) * (1.f - m.p_diffuse);
					else pdf = pdf * m.p_diffuse + blinnPdf(cos_n_h, cos_wo_h, m.gl_exponent) * (1.f - m.p_diffuse);
				}
				return pdf;
			}
			if(use_glossy)
			{
				const V3 h = normalize(wi + wo);
				const float cos_wo_h = dot(wo, h), cos_n_h = dot(n, h);
				if(m.anisotropic) pdf = asAnisoPdf(mk(dot(h, sp.nu), dot(h, sp.nv), cos_n_h), cos_wo_h, m.exp_u, m.exp_v);
				else pdf = blinnPdf(cos_n_h, cos_wo_h, m.gl_exponent);
			}
			return pdf;
		}
		case YAFGPU_MAT_COATED_GLOSSY:
		{	// material_coated_glossy.cc:371-416
			if((dot(sp.ng, wo) * dot(sp.ng, wi)) < 0.f) return 0.f;
			const V3 n = faceForward(sp.ng, sp.n, wo);
			float pdf = 0.f, kr, kt;
			fresnel(wo, n, m.ior, kr, kt);
			float accum_c[3], sum = 0.f;
			accum_c[0] = kr; accum_c[1] = kt * (1.f - m.p_diffuse); accum_c[2] = kt * m.p_diffuse;
			int n_match = 0;
			for(int i = 0; i < sdNBsdf<MT>(m); ++i)
			{
				if((bsdfs & sdCFlags<MT>(m, i)) == sdCFlags<MT>(m, i))
				{
					const float width = accum_c[i];
					sum += width;
					if(i == 1)
					{
						const V3 h = normalize(wi + wo);
						const float cos_wo_h = dot(wo, h), cos_n_h = dot(n, h);
						if(m.anisotropic) pdf += asAnisoPdf(mk(dot(h, sp.nu), dot(h, sp.nv), cos_n_h), cos_wo_h, m.exp_u, m.exp_v) * width;
						else pdf += blinnPdf(cos_n_h, cos_wo_h, m.gl_exponent) * width;
					}
					else if(i == 2) pdf += fabsf(dot(wi, n)) * width;
					++n_match;
				}
			}
			if(!n_match || sum < 0.00001f) return 0.f;
			return pdf / sum;
		}
		default: return 0.f;
	}
}

// Material::sample. Returns the colour; wi, s.pdf, s.sampled and w as the reference leaves them.
template<int MT>
YG_D Col leafSample(const MatDev &m, const SurfPt &sp, V3 wo, V3 &wi, BSample &s, float &w)
{
	switch(mtypeOf<MT>(m))
	{
		case YAFGPU_MAT_SHINYDIFFUSE:
		{	// material_shiny_diffuse.cc:308-408
			float accum_c[4], comp[4];
			const float cos_ng_wo = dot(sp.ng, wo);
			const V3 n = faceForward(sp.ng, sp.n, wo);
			sdComponents<MT>(m, comp);
			const float kr = sdFresnel<MT>(m, wo, n);
			sdAccumulate(comp, accum_c, kr);
			float sum = 0.f, val[4], width[4];
			unsigned choice[4];
			int n_match = 0, pick = -1;
			for(int i = 0; i < sdNBsdf<MT>(m); ++i)
			{
				if((s.flags & sdCFlags<MT>(m, i)) == sdCFlags<MT>(m, i))
				{
					width[n_match] = accum_c[sdCIndex<MT>(m, i)];
					sum += width[n_match];
					choice[n_match] = sdCFlags<MT>(m, i);
					val[n_match] = sum;
					++n_match;
				}
			}
			if(!n_match || sum < 0.00001f) { s.sampled = BsdfNone; s.pdf = 0.f; return mkc(1.f); }
			const float inv_sum = 1.f / sum;
			for(int i = 0; i < n_match; ++i)
			{
				val[i] *= inv_sum;
				width[i] *= inv_sum;
				if((s.s1 <= val[i]) && (pick < 0)) pick = i;
			}
			if(pick < 0) pick = n_match - 1;
			float s_1;
			if(pick > 0) s_1 = (s.s1 - val[pick - 1]) / width[pick];
			else s_1 = s.s1 / width[pick];
			Col scolor = mkc(0.f);
			switch(choice[pick])
			{
				case(BsdfSpecular | BsdfReflect):
					wi = reflectDir(n, wo);
					s.pdf = width[pick];
					scolor = ldc(m.mirror_col) * accum_c[0];
					scolor = scolor * (1.f / fmaxf(fabsf(dot(sp.n, wi)), 1.0e-6f));
					break;
				case(BsdfTransmit | BsdfFilter):
				{
					wi = -wo;
					const float tf = m.transmit_filter_strength;
					scolor = accum_c[1] * (tf * ldc(m.diffuse_col) + mkc(1.f - tf));
					const float cos_n = fabsf(dot(wi, n));
					if(cos_n < 1e-6f) s.pdf = 0.f;
					else s.pdf = width[pick];
					break;
				}
				case(BsdfDiffuse | BsdfTransmit):
					wi = sampleCosHemisphere(-n, sp.nu, sp.nv, s_1, s.s2);
					if(cos_ng_wo * dot(sp.ng, wi) < 0) scolor = accum_c[2] * ldc(m.diffuse_col);
					s.pdf = fabsf(dot(wi, n)) * width[pick];
					break;
				case(BsdfDiffuse | BsdfReflect):
				default:
					wi = sampleCosHemisphere(n, sp.nu, sp.nv, s_1, s.s2);
					if(cos_ng_wo * dot(sp.ng, wi) > 0) scolor = accum_c[3] * ldc(m.diffuse_col);
					if(m.use_oren_nayar) scolor = scolor * ((MT < 0 && m.oren_tex) ? orenNayarTex(m.oren_sigma, wo, wi, n) : orenNayar(m.oren_a, m.oren_b, wo, wi, n));
					s.pdf = fabsf(dot(wi, n)) * width[pick];
					break;
			}
			s.sampled = choice[pick];
			w = (fabsf(dot(wi, sp.n))) / (s.pdf * 0.99f + 0.01f);
			// getAlpha :569-597
			float alpha = 1.f;
			if(!YG_SDL(MT) && m.is_transparent) alpha = 1.f - (1.f - comp[0] * kr) * comp[1];
			w = w * (alpha) + 1.f * (1.f - alpha);
			return scolor;
		}
		case YAFGPU_MAT_GLASS:
		{	// material_glass.cc:64-216
			if(!(s.flags & BsdfSpecular) && !((s.flags & m.bsdf_flags & BsdfDispersive) && sp.chromatic)) { s.pdf = 0.f; return mkc(0.f); }
			const bool outside = dot(sp.ng, wo) > 0;
			const V3 n = glassNormal(sp, wo, outside);
			s.pdf = 1.f;
			V3 refdir;
			const float cur_ior = YG_IOR(m, sp);
			const bool spec_refl = (s.flags & (BsdfSpecular | BsdfReflect)) == (BsdfSpecular | BsdfReflect);
			if(m.disperse && sp.chromatic)
			{	// :83-138 the dispersion has to be sampled: the refracted direction belongs to one wavelength
				if(refract(n, wo, refdir, cur_ior))
				{
					float kr, kt;
					fresnel(wo, n, cur_ior, kr, kt);
					const float p_kr = (float)(0.01 + 0.99 * (double)kr), p_kt = (float)(0.01 + 0.99 * (double)kt);
					if(!(s.flags & BsdfSpecular) || s.s1 < p_kt)
					{
						wi = refdir; s.pdf = spec_refl ? p_kt : 1.f; s.sampled = BsdfDispersive | BsdfTransmit; w = 1.f;
						return ldc(m.filter_col);
					}
					else if(spec_refl)
					{
						wi = reflect(wo, n); s.pdf = p_kr; s.sampled = BsdfSpecular | BsdfReflect; w = 1.f;
						return ldc(m.spec_refl_col);
					}
				}
				else if(spec_refl)
				{
					wi = reflect(wo, n); s.sampled = BsdfSpecular | BsdfReflect; w = 1.f;
					return mkc(1.f);
				}
				s.pdf = 0.f;
				return mkc(0.f);
			}
			if(refract(n, wo, refdir, cur_ior))
			{
				float kr, kt;
				fresnel(wo, n, cur_ior, kr, kt);
				const float p_kr = (float)(0.01 + 0.99 * (double)kr), p_kt = (float)(0.01 + 0.99 * (double)kt);
				if(s.s1 < p_kt && ((s.flags & m.tm_flags) == m.tm_flags))
				{
					wi = refdir; s.pdf = p_kt; s.sampled = m.tm_flags; w = 1.f;
					return ldc(m.filter_col);
				}
				else if(spec_refl)
				{
					wi = reflect(wo, n); s.pdf = p_kr; s.sampled = BsdfSpecular | BsdfReflect; w = 1.f;
					return ldc(m.spec_refl_col);
				}
			}
			else if(spec_refl)
			{
				wi = reflect(wo, n); s.sampled = BsdfSpecular | BsdfReflect; w = 1.f;
				return mkc(1.f);
			}
			s.pdf = 0.f;
			return mkc(0.f);
		}
		case YAFGPU_MAT_MIRROR: // material_glass.cc:466-472
			wi = reflectDir(sp.n, wo);
			s.sampled = BsdfSpecular | BsdfReflect;
			w = 1.f;
			return ldc(m.simple_col) * (1.f / fabsf(dot(sp.n, wi)));
		case YAFGPU_MAT_GLOSSY:
		{	// material_glossy.cc:176-357
			const float cos_ng_wo = dot(sp.ng, wo);
			const V3 n = faceForward(sp.ng, sp.n, wo);
			s.pdf = 0.f;
			float wi_n = 0.f;
			const float wo_n = fabsf(dot(wo, n));
			float cos_wo_h = 0.f;
			Col scolor = mkc(0.f);
			float s_1 = s.s1;
			const float cur_p_diffuse = m.p_diffuse;
			const bool use_glossy = m.as_diffuse ? (s.flags & BsdfDiffuse) : (s.flags & BsdfGlossy);
			const bool use_diffuse = m.with_diffuse && (s.flags & BsdfDiffuse);
			float glossy = 0.f;
			if(use_diffuse)
			{
				const float s_p_diffuse = use_glossy ? cur_p_diffuse : 1.f;
				if(s_1 < s_p_diffuse)
				{
					s_1 /= s_p_diffuse;
					wi = sampleCosHemisphere(n, sp.nu, sp.nv, s_1, s.s2);
					if(dot(sp.ng, wi) * cos_ng_wo < 0.f) return scolor;
					wi_n = fabsf(dot(wi, n));
					s.pdf = wi_n;
					if(use_glossy)
					{
						const V3 h = normalize(wi + wo);
						cos_wo_h = dot(wo, h);
						const float cos_wi_h = fabsf(dot(wi, h));
						const float cos_n_h = dot(n, h);
						if(m.anisotropic)
						{
							const V3 hs = mk(dot(h, sp.nu), dot(h, sp.nv), cos_n_h);
							s.pdf = s.pdf * cur_p_diffuse + asAnisoPdf(hs, cos_wo_h, m.exp_u, m.exp_v) * (1.f - cur_p_diffuse);
							glossy = divAsDivisor(asAnisoD(hs, m.exp_u, m.exp_v) * schlickFresnel(cos_wi_h, m.gl_glossy), cos_wi_h, wo_n, wi_n);
						}
						else
						{
							s.pdf = s.pdf * cur_p_diffuse + blinnPdf(cos_n_h, cos_wo_h, m.gl_exponent) * (1.f - cur_p_diffuse);
							glossy = divAsDivisor(blinnD(cos_n_h, m.gl_exponent) * schlickFresnel(cos_wi_h, m.gl_glossy), cos_wi_h, wo_n, wi_n);
						}
					}
					s.sampled = BsdfDiffuse | BsdfReflect;
					if(!(s.flags & BsdfReflect)) return mkc(0.f);
					scolor = glossy * ldc(m.gloss_col);
					{
						Col add_col = diffuseReflect(wi_n, wo_n, m.gl_glossy, m.gl_diffuse, ldc(m.diff_col));
						if(MT < 0) add_col = add_col * m.diffuse_refl_mul; // "diffuse_refl_shader"; 1 when absent
						if(m.gl_oren_nayar) add_col = add_col * ((MT < 0 && m.oren_tex) ? orenNayarTex(m.oren_sigma, wi, wo, n) : orenNayar(m.gl_oren_a, m.gl_oren_b, wi, wo, n));
						scolor = scolor + add_col;
					}
					w = wi_n / (s.pdf * 0.99f + 0.01f);
					return scolor;
				}
				s_1 -= cur_p_diffuse;
				s_1 /= (1.f - cur_p_diffuse);
			}
			if(use_glossy)
			{
				const V3 Hs = m.anisotropic ? asAnisoSample(s_1, s.s2, m.exp_u, m.exp_v) : blinnSample(s_1, s.s2, m.gl_exponent);
				V3 h = (Hs.x * sp.nu + Hs.y * sp.nv) + Hs.z * n;
				cos_wo_h = dot(wo, h);
				if(cos_wo_h < 0.f) { h = reflect(h, n); cos_wo_h = dot(wo, h); }
				wi = reflectDir(h, wo);
				if(cos_ng_wo * dot(sp.ng, wi) < 0.f) return mkc(0.f);
				wi_n = fabsf(dot(wi, n));
				if(m.anisotropic)
				{
					s.pdf = asAnisoPdf(Hs, cos_wo_h, m.exp_u, m.exp_v);
					glossy = divAsDivisor(asAnisoD(Hs, m.exp_u, m.exp_v) * schlickFresnel(cos_wo_h, m.gl_glossy), cos_wo_h, wo_n, wi_n);
				}
				else
				{
					const float cos_hn = dot(h, n);
					s.pdf = blinnPdf(cos_hn, cos_wo_h, m.gl_exponent);
					glossy = divAsDivisor(blinnD(cos_hn, m.gl_exponent) * schlickFresnel(cos_wo_h, m.gl_glossy), cos_wo_h, wo_n, wi_n);
				}
				scolor = glossy * ldc(m.gloss_col);
				s.sampled = m.as_diffuse ? (BsdfDiffuse | BsdfReflect) : (BsdfGlossy | BsdfReflect);
			}
			if(use_diffuse)
			{
				Col add_col = diffuseReflect(wi_n, wo_n, m.gl_glossy, m.gl_diffuse, ldc(m.diff_col));
				if(MT < 0) add_col = add_col * m.diffuse_refl_mul; // "diffuse_refl_shader"; 1 when absent
				if(m.gl_oren_nayar) add_col = add_col * ((MT < 0 && m.oren_tex) ? orenNayarTex(m.oren_sigma, wi, wo, n) : orenNayar(m.gl_oren_a, m.gl_oren_b, wi, wo, n));
				s.pdf = wi_n * cur_p_diffuse + s.pdf * (1.f - cur_p_diffuse);
				scolor = scolor + add_col;
			}
			w = wi_n / (s.pdf * 0.99f + 0.01f);
			return scolor;
		}
		case YAFGPU_MAT_COATED_GLOSSY:
		{	// material_coated_glossy.cc:187-369
			const float cos_ng_wo = dot(sp.ng, wo);
			const V3 n = faceForward(sp.ng, sp.n, wo);
			V3 hs = mk(0.f);
			s.pdf = 0.f;
			float kr, kt, wi_n = 0.f, wo_n = 0.f;
			fresnel(wo, n, m.ior, kr, kt);
			bool use[3] = {false, false, false};
			float sum = 0.f, accum_c[3], val[3], width[3];
			int c_index[3], rc_index[3] = {0, 0, 0};
			accum_c[0] = kr; accum_c[1] = kt * (1.f - m.p_diffuse); accum_c[2] = kt * m.p_diffuse;
			int n_match = 0, pick = -1;
			for(int i = 0; i < sdNBsdf<MT>(m); ++i)
			{
				if((s.flags & sdCFlags<MT>(m, i)) == sdCFlags<MT>(m, i))
				{
					use[i] = true;
					width[n_match] = accum_c[i];
					c_index[n_match] = i;
					rc_index[i] = n_match;
					sum += width[n_match];
					val[n_match] = sum;
					++n_match;
				}
			}
			if(!n_match || sum < 0.00001f) { wi = reflectDir(n, wo); return mkc(0.f); }
			else if(n_match == 1) { pick = 0; width[0] = 1.f; }
			else
			{
				const float inv_sum = 1.f / sum;
				for(int i = 0; i < n_match; ++i)
				{
					val[i] *= inv_sum;
					width[i] *= inv_sum;
					if((s.s1 <= val[i]) && (pick < 0)) pick = i;
				}
			}
			if(pick < 0) pick = n_match - 1;
			float s_1;
			if(pick > 0) s_1 = (s.s1 - val[pick - 1]) / width[pick];
			else s_1 = s.s1 / width[pick];
			Col scolor = mkc(0.f);
			switch(c_index[pick])
			{
				case 0:
					wi = reflectDir(n, wo);
					scolor = ldc(m.gl_mirror_col) * kr;
					scolor = scolor * m.gl_mirror_strength;
					s.pdf = width[pick];
					break;
				case 1:
					hs = m.anisotropic ? asAnisoSample(s_1, s.s2, m.exp_u, m.exp_v) : blinnSample(s_1, s.s2, m.gl_exponent);
					break;
				default:
					wi = sampleCosHemisphere(n, sp.nu, sp.nv, s_1, s.s2);
					if(cos_ng_wo * dot(sp.ng, wi) < 0) return mkc(0.f);
			}
			wi_n = fabsf(dot(wi, n)); // for the glossy pick wi is not set yet in the reference either (recomputed below)
			wo_n = fabsf(dot(wo, n));
			if(c_index[pick] != 0)
			{
				if(use[1])
				{
					float glossy, cos_wo_h;
					V3 h;
					if(c_index[pick] != 1)
					{
						h = normalize(wi + wo);
						hs = mk(dot(h, sp.nu), dot(h, sp.nv), dot(h, n));
						cos_wo_h = dot(wo, h);
					}
					else
					{
						h = (hs.x * sp.nu + hs.y * sp.nv) + hs.z * n;
						cos_wo_h = dot(wo, h);
						if(cos_wo_h < 0.f) { h = reflect(h, n); cos_wo_h = dot(wo, h); }
						wi = reflectDir(h, wo);
						if(cos_ng_wo * dot(sp.ng, wi) < 0) return mkc(0.f);
					}
					wi_n = fabsf(dot(wi, n));
					if(m.anisotropic)
					{
						s.pdf += asAnisoPdf(hs, cos_wo_h, m.exp_u, m.exp_v) * width[rc_index[1]];
						glossy = divAsDivisor(asAnisoD(hs, m.exp_u, m.exp_v) * schlickFresnel(cos_wo_h, m.gl_glossy), cos_wo_h, wo_n, wi_n);
					}
					else
					{
						const float cos_hn = dot(h, n);
						s.pdf += blinnPdf(cos_hn, cos_wo_h, m.gl_exponent) * width[rc_index[1]];
						glossy = divAsDivisor(blinnD(cos_hn, m.gl_exponent) * schlickFresnel(cos_wo_h, m.gl_glossy), cos_wo_h, wo_n, wi_n);
					}
					scolor = (glossy * kt) * ldc(m.gloss_col);
				}
				if(use[2])
				{
					Col add_col = kt * diffuseReflect(wi_n, wo_n, m.gl_glossy, m.gl_diffuse, ldc(m.diff_col));
					if(MT < 0) add_col = add_col * m.diffuse_refl_mul; // "diffuse_refl_shader"; 1 when absent
					if(m.gl_oren_nayar) add_col = add_col * ((MT < 0 && m.oren_tex) ? orenNayarTex(m.oren_sigma, wi, wo, n) : orenNayar(m.gl_oren_a, m.gl_oren_b, wi, wo, n));
					scolor = scolor + add_col;
					s.pdf += wi_n * width[rc_index[2]];
				}
				w = wi_n / (s.pdf * 0.99f + 0.01f);
			}
			else w = 1.f;
			s.sampled = m.c_flags[c_index[pick]];
			return scolor;
		}
		case YAFGPU_MAT_ROUGH_GLASS:
		{	// material_rough_glass.cc:61-172
			const V3 n = faceForward(sp.ng, sp.n, wo);
			const bool outside = dot(sp.ng, wo) > 0.f;
			s.pdf = 1.f;
			const float alpha_2 = m.alpha_rough * m.alpha_rough;
			V3 h = ggxSample(alpha_2, s.s1, s.s2);
			h = normalize((h.x * sp.nu + h.y * sp.nv) + h.z * n);
			const float cur_ior = YG_IOR(m, sp);
			float glossy, glossy_d = 0.f, glossy_g = 0.f, wi_n, wi_h, jacobian;
			const float cos_theta = dot(h, n);
			const float cos_theta_2 = cos_theta * cos_theta;
			const float tan_theta_2 = (1.f - cos_theta_2) / fmaxf(1.0e-8f, cos_theta_2);
			if(cos_theta > 0.f) glossy_d = ggxD(alpha_2, cos_theta_2, tan_theta_2);
			const float wo_h = dot(wo, h), wo_n = dot(wo, n);
			float kr, kt;
			Col ret = mkc(0.f);
			if(refractMicrofacet(outside ? 1.f / cur_ior : cur_ior, wo, wi, h, wo_h, kr, kt))
			{
				if(s.s1 < kt && (s.flags & BsdfTransmit))
				{
					wi_n = dot(wi, n); wi_h = dot(wi, h);
					if((wi_h * wi_n) > 0.f && (wo_h * wo_n) > 0.f) glossy_g = ggxG(alpha_2, wi_n, wo_n);
					float ior_wi = 1.f, ior_wo = 1.f;
					if(outside) ior_wi = cur_ior; else ior_wo = cur_ior;
					const float ht = ior_wo * wo_h + ior_wi * wi_h;
					jacobian = (ior_wi * ior_wi) / fmaxf(1.0e-8f, ht * ht);
					glossy = fabsf((wo_h * wi_h) / (wi_n * wo_n)) * kt * glossy_g * glossy_d * jacobian;
					s.pdf = glossy_d * cos_theta * (jacobian * fabsf(wi_h));
					s.sampled = ((m.disperse && sp.chromatic) ? BsdfDispersive : BsdfGlossy) | BsdfTransmit;
					ret = glossy * ldc(m.filter_col);
					w = fabsf(wi_n) / fmaxf(0.1f, s.pdf);
				}
				else if(s.flags & BsdfReflect)
				{
					wi = reflectMicrofacet(wo, h);
					wi_n = dot(wi, n); wi_h = dot(wi, h);
					glossy_g = ggxG(alpha_2, wi_n, wo_n);
					jacobian = 1.f / fmaxf(1.0e-8f, (4.f * fabsf(wi_h)));
					glossy = (kr * glossy_g * glossy_d) / fmaxf(1.0e-8f, (4.f * fabsf(wo_n * wi_n)));
					s.pdf = glossy_d * cos_theta * jacobian;
					s.sampled = BsdfGlossy | BsdfReflect;
					ret = glossy * ldc(m.spec_refl_col);
					w = fabsf(wi_n) / fmaxf(0.1f, s.pdf);
				}
			}
			else
			{
				wi = reflect(wo, h);
				s.sampled = BsdfGlossy | BsdfReflect;
				ret = mkc(1.f);
				w = 1.f;
			}
			return ret;
		}
		default: // null, light_mat: material_glass.cc:495-500, material_simple.cc:43-48
			s.pdf = 0.f; w = 0.f;
			return mkc(0.f);
	}
}

// two-direction sample (only RoughGlass implements it; the base class returns black and touches nothing, material.h:131)
template<int MT>
YG_D Col leafSample2(const MatDev &m, const SurfPt &sp, V3 wo, V3 *dir, Col &tcol, BSample &s, float *w)
{
	if(mtypeOf<MT>(m) != YAFGPU_MAT_ROUGH_GLASS) return mkc(0.f);
	// material_rough_glass.cc:174-291
	const V3 n = faceForward(sp.ng, sp.n, wo);
	const bool outside = dot(sp.ng, wo) > 0.f;
	s.pdf = 1.f;
	const float alpha_2 = m.alpha_rough * m.alpha_rough;
	V3 h = ggxSample(alpha_2, s.s1, s.s2), wi;
	h = normalize((h.x * sp.nu + h.y * sp.nv) + h.z * n);
	const float cur_ior = YG_IOR(m, sp);
	float glossy, glossy_d = 0.f, glossy_g = 0.f, wi_n, wi_h, jacobian;
	const float cos_theta = dot(h, n);
	const float cos_theta_2 = cos_theta * cos_theta;
	const float tan_theta_2 = (1.f - cos_theta_2) / fmaxf(1.0e-8f, cos_theta_2);
	if(cos_theta > 0.f) glossy_d = ggxD(alpha_2, cos_theta_2, tan_theta_2);
	const float wo_h = dot(wo, h), wo_n = dot(wo, n);
	float kr, kt;
	Col ret = mkc(0.f);
	s.sampled = 0;
	if(refractMicrofacet(outside ? 1.f / cur_ior : cur_ior, wo, wi, h, wo_h, kr, kt))
	{
		if(s.flags & BsdfTransmit)
		{
			wi_n = dot(wi, n); wi_h = dot(wi, h);
			if((wi_h * wi_n) > 0.f && (wo_h * wo_n) > 0.f) glossy_g = ggxG(alpha_2, wi_n, wo_n);
			float ior_wi = 1.f, ior_wo = 1.f;
			if(outside) ior_wi = cur_ior; else ior_wo = cur_ior;
			const float ht = ior_wo * wo_h + ior_wi * wi_h;
			jacobian = (ior_wi * ior_wi) / fmaxf(1.0e-8f, ht * ht);
			glossy = fabsf((wo_h * wi_h) / (wi_n * wo_n)) * kt * glossy_g * glossy_d * jacobian;
			s.pdf = glossy_d * cos_theta * (jacobian * fabsf(wi_h));
			s.sampled = ((m.disperse && sp.chromatic) ? BsdfDispersive : BsdfGlossy) | BsdfTransmit;
			ret = glossy * ldc(m.filter_col);
			w[0] = fabsf(wi_n) / fmaxf(0.1f, s.pdf);
			dir[0] = wi;
#ifdef YG_NAN_TRAP
			if(!isfinite(glossy) || !isfinite(w[0])) printf("rough glass transmit lobe: glossy %g w %g pdf %g | wo_h %g wi_h %g wi_n %g wo_n %g kt %g G %g D %g jac %g ht %g cos_theta %g tan2 %g s1 %g s2 %g h (%g %g %g) n (%g %g %g) nu (%g %g %g) nv (%g %g %g)\n",
				glossy, w[0], s.pdf, wo_h, wi_h, wi_n, wo_n, kt, glossy_g, glossy_d, jacobian, ht, cos_theta, tan_theta_2, s.s1, s.s2, h.x, h.y, h.z, n.x, n.y, n.z, sp.nu.x, sp.nu.y, sp.nu.z, sp.nv.x, sp.nv.y, sp.nv.z);
#endif
		}
		if(s.flags & BsdfReflect)
		{
			wi = reflectMicrofacet(wo, h);
			wi_n = dot(wi, n); wi_h = dot(wi, h);
			glossy_g = ggxG(alpha_2, wi_n, wo_n);
			jacobian = 1.f / fmaxf(1.0e-8f, (4.f * fabsf(wi_h)));
			glossy = (kr * glossy_g * glossy_d) / fmaxf(1.0e-8f, (4.f * fabsf(wo_n * wi_n)));
			s.pdf = glossy_d * cos_theta * jacobian;
			s.sampled |= BsdfGlossy | BsdfReflect;
			tcol = glossy * ldc(m.spec_refl_col);
			w[1] = fabsf(wi_n) / fmaxf(0.1f, s.pdf);
			dir[1] = wi;
		}
	}
	else
	{
		wi = reflect(wo, h);
		s.sampled |= BsdfGlossy | BsdfReflect;
		dir[0] = wi;
		ret = mkc(1.f);
		w[0] = 1.f;
	}
	return ret;
}

template<int MT>
YG_D void leafSpecular(const MatDev &m, const SurfPt &sp, V3 wo, int raylevel, bool &refl, bool &refr, V3 *dir, Col *col)
{
	switch(mtypeOf<MT>(m))
	{
		case YAFGPU_MAT_SHINYDIFFUSE:
		{	// material_shiny_diffuse.cc:474-528
			const bool backface = dot(wo, sp.ng) < 0.f;
			const V3 n = backface ? -sp.n : sp.n;
			const V3 ng = backface ? -sp.ng : sp.ng;
			float comp[4]; sdComponents<MT>(m, comp);
			const float kr = sdFresnel<MT>(m, wo, n);
			if(!YG_SDL(MT) && m.is_transparent)
			{
				refr = true;
				dir[1] = -wo;
				const float tf = m.transmit_filter_strength;
				const Col tcol = tf * ldc(m.diffuse_col) + mkc(1.f - tf);
				col[1] = ((1.f - comp[0] * kr) * comp[1]) * tcol;
			}
			else refr = false;
			if(!YG_SDL(MT) && m.is_mirror)
			{
				refl = true;
				dir[0] = reflect(wo, n);
				const float cos_wi_ng = dot(dir[0], ng);
				if((double)cos_wi_ng < 0.01)
				{
					dir[0] = dir[0] + (float)(0.01 - (double)cos_wi_ng) * ng;
					dir[0] = normalize(dir[0]);
				}
				col[0] = ldc(m.mirror_col) * (comp[0] * kr);
			}
			else refl = false;
			return;
		}
		case YAFGPU_MAT_GLASS:
		{	// material_glass.cc:242-309
			const bool outside = dot(sp.ng, wo) > 0;
			const V3 n = glassNormal(sp, wo, outside);
			V3 refdir;
			const float cur_ior = YG_IOR(m, sp);
			if(refract(n, wo, refdir, cur_ior))
			{
				float kr, kt;
				fresnel(wo, n, cur_ior, kr, kt);
				if(!sp.chromatic || !m.disperse)
				{
					col[1] = kt * ldc(m.filter_col);
					dir[1] = refdir;
					refr = true;
				}
				else refr = false; // the refraction is sampled per wavelength by the dispersive branch of recursiveRaytrace, not specular here
				if(outside || raylevel < 3)
				{
					dir[0] = reflect(wo, n);
					col[0] = ldc(m.spec_refl_col) * kr;
					refl = true;
				}
				else refl = false;
			}
			else
			{
				col[0] = ldc(m.spec_refl_col);
				dir[0] = reflect(wo, n);
				refl = true; refr = false;
			}
			return;
		}
		case YAFGPU_MAT_MIRROR:
		{	// material_glass.cc:474-483
			col[0] = ldc(m.simple_col); col[1] = mkc(1.f);
			const V3 n = faceForward(sp.ng, sp.n, wo);
			dir[0] = reflectDir(n, wo);
			refl = true; refr = false;
			return;
		}
		case YAFGPU_MAT_COATED_GLOSSY:
		{	// material_coated_glossy.cc:418-463
			const bool outside = dot(sp.ng, wo) >= 0;
			const V3 n = glassNormal(sp, wo, outside);
			const V3 ng = outside ? sp.ng : -sp.ng;
			float kr, kt;
			fresnel(wo, n, m.ior, kr, kt);
			refr = false;
			if(raylevel > 5) return;
			dir[0] = reflect(wo, n);
			col[0] = (ldc(m.gl_mirror_col) * kr) * m.gl_mirror_strength;
			const float cos_wi_ng = dot(dir[0], ng);
			if((double)cos_wi_ng < 0.01)
			{
				dir[0] = dir[0] + (float)(0.01 - (double)cos_wi_ng) * ng;
				dir[0] = normalize(dir[0]);
			}
			refl = true;
			return;
		}
		default: refl = false; refr = false; return; // Material::getSpecular default material.h:151-153
	}
}

template<int MT>
YG_D Col leafEmit(const MatDev &m, const SurfPt &sp, V3 wo, bool include_lights)
{
	if(mtypeOf<MT>(m) == YAFGPU_MAT_SHINYDIFFUSE) return ldc(m.emit_col); // :295-306
	if(mtypeOf<MT>(m) == YAFGPU_MAT_LIGHT)
	{	// material_simple.cc:50-58
		if(!include_lights) return mkc(0.f);
		if(m.double_sided) return ldc(m.simple_col);
		return (dot(wo, sp.n) > 0) ? ldc(m.simple_col) : mkc(0.f);
	}
	return mkc(0.f);
}

// Material::getTransparency (only called when isTransparent())
template<int MT>
YG_D Col leafTransparency(const MatDev &m, const SurfPt &sp, V3 wo)
{
	if(mtypeOf<MT>(m) == YAFGPU_MAT_SHINYDIFFUSE)
	{	// :530-567
		if(YG_SDL(MT) || !m.is_transparent) return mkc(0.f);
		float accum = 1.f;
		const V3 n = faceForward(sp.ng, sp.n, wo);
		const float kr = sdFresnel<MT>(m, wo, n);
		if(m.is_mirror) accum = 1.f - kr * m.mirror_strength;
		accum *= m.transparency_strength * accum;
		const float tf = m.transmit_filter_strength;
		return accum * (tf * ldc(m.diffuse_col) + mkc(1.f - tf));
	}
	if(mtypeOf<MT>(m) == YAFGPU_MAT_GLASS || mtypeOf<MT>(m) == YAFGPU_MAT_ROUGH_GLASS)
	{	// material_glass.cc:218-229, material_rough_glass.cc:293-303
		const V3 n = faceForward(sp.ng, sp.n, wo);
		float kr, kt;
		fresnel(wo, n, m.ior, kr, kt);
		return kt * ldc(m.filter_col);
	}
	return mkc(0.f);
}

// Material::getAlpha (material.h:163 = 1; the wireframe term is refused at upload): what "bg_transp_refract" folds along the refraction chain
template<int MT>
YG_D float leafAlpha(const MatDev &m, const SurfPt &sp, V3 wo)
{
	if(mtypeOf<MT>(m) == YAFGPU_MAT_SHINYDIFFUSE)
	{	// material_shiny_diffuse.cc:569-597
		if(YG_SDL(MT) || !m.is_transparent) return 1.f;
		const V3 n = faceForward(sp.ng, sp.n, wo);
		float comp[4]; sdComponents<MT>(m, comp);
		const float kr = sdFresnel<MT>(m, wo, n);
		const float refl = (1.f - comp[0] * kr) * comp[1];
		return 1.f - refl;
	}
	if(mtypeOf<MT>(m) == YAFGPU_MAT_GLASS)
	{	// material_glass.cc:231-240: `1.0 - energy` is a double subtraction
		const Col t = leafTransparency<MT>(m, sp, wo);
		float alpha = (float)(1.0 - (double)((t.r + t.g + t.b) * 0.333333f));
		if(alpha < 0.0f) alpha = 0.0f;
		return alpha;
	}
	if(mtypeOf<MT>(m) == YAFGPU_MAT_ROUGH_GLASS)
	{	// material_rough_glass.cc:306-315
		const Col t = leafTransparency<MT>(m, sp, wo);
		return fmaxf(0.f, fminf(1.f, 1.f - (t.r + t.g + t.b) * 0.333333f));
	}
	return 1.f;
}

// BlendMaterial::initBsdf replaces the surface point by blendSurfacePoints__(sp, sp, ival) (material_blend.cc:77-104,
// surface.cc:137-151): lerp__(a, a, alpha) = a * alpha + a * (1 - alpha) in float, which is not exactly `a`.
YG_D V3 blendLerp(V3 a, float al, float ial) { return a * al + a * ial; }
template<int MT>
YG_D void matInitBsdf(const MatDev &m, SurfPt &sp)
{
	if(mtypeOf<MT>(m) != YAFGPU_MAT_BLEND) return;
	const float ival = fminf(1.f, fmaxf(0.f, 1.f - m.blend_val));
	const float al = ival, ial = (float)(1.0 - (double)ival);
	sp.n = blendLerp(sp.n, al, ial); sp.nu = blendLerp(sp.nu, al, ial); sp.nv = blendLerp(sp.nv, al, ial);
}

// ---- top level: BlendMaterial dispatch (children are leaf materials; nesting is refused at upload) ----------------------
// The children of a blend are reached through ONE out-of-line copy of each generic leaf function (type read from the record),
// so the blend kernel's code stays small; every other kernel inlines exactly its own material's code.
YG_DN Col leafEvalG(const MatDev &m, const SurfPt &sp, V3 wo, V3 wl, unsigned bsdfs) { return leafEval<-1>(m, sp, wo, wl, bsdfs); }
YG_DN float leafPdfG(const MatDev &m, const SurfPt &sp, V3 wo, V3 wi, unsigned bsdfs) { return leafPdf<-1>(m, sp, wo, wi, bsdfs); }
YG_DN Col leafSampleG(const MatDev &m, const SurfPt &sp, V3 wo, V3 &wi, BSample &s, float &w) { return leafSample<-1>(m, sp, wo, wi, s, w); }
YG_DN Col leafSample2G(const MatDev &m, const SurfPt &sp, V3 wo, V3 *dir, Col &tcol, BSample &s, float *w) { return leafSample2<-1>(m, sp, wo, dir, tcol, s, w); }
YG_DN void leafSpecularG(const MatDev &m, const SurfPt &sp, V3 wo, int raylevel, bool &refl, bool &refr, V3 *dir, Col *col) { leafSpecular<-1>(m, sp, wo, raylevel, refl, refr, dir, col); }
YG_DN Col leafEmitG(const MatDev &m, const SurfPt &sp, V3 wo, bool include_lights) { return leafEmit<-1>(m, sp, wo, include_lights); }
YG_DN Col leafTransparencyG(const MatDev &m, const SurfPt &sp, V3 wo) { return leafTransparency<-1>(m, sp, wo); }
YG_DN float leafAlphaG(const MatDev &m, const SurfPt &sp, V3 wo) { return leafAlpha<-1>(m, sp, wo); }

#define YG_NOT_BLEND(MT, m) ((MT) >= 0 ? (MT) != YAFGPU_MAT_BLEND : (m).type != YAFGPU_MAT_BLEND)

template<int MT>
YG_D Col matEval(const SceneDev &sc, const MatDev &m, const SurfPt &sp, V3 wo, V3 wl, unsigned bsdfs)
{
	if(YG_NOT_BLEND(MT, m)) return leafEval<MT>(m, sp, wo, wl, bsdfs);
	const float val = m.blend_val, ival = fminf(1.f, fmaxf(0.f, 1.f - val));
	const Col c1 = leafEvalG(sc.materials[m.mat1], sp, wo, wl, bsdfs), c2 = leafEvalG(sc.materials[m.mat2], sp, wo, wl, bsdfs);
	return c1 * ival + c2 * val;
}
template<int MT>
YG_D float matPdf(const SceneDev &sc, const MatDev &m, const SurfPt &sp, V3 wo, V3 wi, unsigned bsdfs)
{
	if(YG_NOT_BLEND(MT, m)) return leafPdf<MT>(m, sp, wo, wi, bsdfs);
	const float val = m.blend_val, ival = fminf(1.f, fmaxf(0.f, 1.f - val));
	return leafPdfG(sc.materials[m.mat1], sp, wo, wi, bsdfs) * ival + leafPdfG(sc.materials[m.mat2], sp, wo, wi, bsdfs) * val;
}
template<int MT>
YG_D Col matSample(const SceneDev &sc, const MatDev &m, const SurfPt &sp, V3 wo, V3 &wi, BSample &s, float &w)
{
	if(YG_NOT_BLEND(MT, m)) return leafSample<MT>(m, sp, wo, wi, s, w);
	// material_blend.cc:131-214
	const float val = m.blend_val, ival = fminf(1.f, fmaxf(0.f, 1.f - val));
	const MatDev &m1 = sc.materials[m.mat1], &m2 = sc.materials[m.mat2];
	bool s1d = false, s2d = false;
	Col col1 = mkc(0.f), col2 = mkc(0.f);
	BSample s1 = s, s2 = s;
	V3 wi_1 = mk(0.f), wi_2 = mk(0.f);
	float w_1 = 0.f, w_2 = 0.f;
	s2.pdf = s1.pdf = s.pdf = 0.f;
	if(s.flags & m1.bsdf_flags) { col1 = leafSampleG(m1, sp, wo, wi_1, s1, w_1); s1d = true; }
	if(s.flags & m2.bsdf_flags) { col2 = leafSampleG(m2, sp, wo, wi_2, s2, w_2); s2d = true; }
	if(s1d && s2d)
	{
		wi = normalize(wi_2 + wi_1);
		s.pdf = s1.pdf * ival + s2.pdf * val;
		s.sampled = s1.sampled | s2.sampled;
		col1 = (col1 * w_1) * ival + (col2 * w_2) * val;
		w = 1.f;
	}
	else if(s1d) { wi = wi_1; s.pdf = s1.pdf; s.sampled = s1.sampled; w = w_1; }
	else if(s2d) { wi = wi_2; s.pdf = s2.pdf; s.sampled = s2.sampled; col1 = col2; w = w_2; }
	return col1;
}
template<int MT>
YG_D Col matSample2(const SceneDev &sc, const MatDev &m, const SurfPt &sp, V3 wo, V3 *dir, Col &tcol, BSample &s, float *w)
{
	if(YG_NOT_BLEND(MT, m)) return leafSample2<MT>(m, sp, wo, dir, tcol, s, w);
	// material_blend.cc:216-237
	const float val = m.blend_val, ival = fminf(1.f, fmaxf(0.f, 1.f - val));
	if(val <= 0.f) return leafSample2G(sc.materials[m.mat1], sp, wo, dir, tcol, s, w);
	if(val >= 1.f) return leafSample2G(sc.materials[m.mat2], sp, wo, dir, tcol, s, w);
	const Col a = leafSample2G(sc.materials[m.mat1], sp, wo, dir, tcol, s, w);
	const Col b = leafSample2G(sc.materials[m.mat2], sp, wo, dir, tcol, s, w);
	return a * ival + b * val;
}
template<int MT>
YG_D void matSpecular(const SceneDev &sc, const MatDev &m, const SurfPt &sp, V3 wo, int raylevel, bool &refl, bool &refr, V3 *dir, Col *col)
{
	if(YG_NOT_BLEND(MT, m)) { leafSpecular<MT>(m, sp, wo, raylevel, refl, refr, dir, col); return; }
	// material_blend.cc:259-325
	const float val = m.blend_val, ival = fminf(1.f, fmaxf(0.f, 1.f - val));
	refl = false; refr = false;
	bool r1 = false, t1 = false;
	V3 d1[2] = {mk(0.f), mk(0.f)};
	Col c1[2] = {mkc(0.f), mkc(0.f)};
	leafSpecularG(sc.materials[m.mat1], sp, wo, raylevel, r1, t1, d1, c1);
	leafSpecularG(sc.materials[m.mat2], sp, wo, raylevel, refl, refr, dir, col);
	if(refl && r1) { col[0] = c1[0] * ival + col[0] * val; dir[0] = normalize(dir[0] + d1[0]); }
	else if(r1) { col[0] = c1[0] * ival; dir[0] = d1[0]; }
	else col[0] = col[0] * val;
	if(refr && t1) { col[1] = c1[1] * ival + col[1] * val; dir[1] = normalize(dir[1] + d1[1]); }
	else if(t1) { col[1] = c1[1] * ival; dir[1] = d1[1]; }
	else col[1] = col[1] * val;
	refr = refr || t1;
	refl = refl || r1;
}
template<int MT>
YG_D Col matEmit(const SceneDev &sc, const MatDev &m, const SurfPt &sp, V3 wo, bool include_lights)
{
	if(YG_NOT_BLEND(MT, m)) return leafEmit<MT>(m, sp, wo, include_lights);
	const float val = m.blend_val, ival = fminf(1.f, fmaxf(0.f, 1.f - val));
	return leafEmitG(sc.materials[m.mat1], sp, wo, include_lights) * ival + leafEmitG(sc.materials[m.mat2], sp, wo, include_lights) * val;
}
// BlendMaterial::getAlpha material_blend.cc:363-395 (the smaller of the two children's, only when one of them isTransparent()); a mask_mat has
// been resolved to the chosen child by evalMaterialNodes (material_mask.cc:122-131)
template<int MT>
YG_D float matAlpha(const SceneDev &sc, const MatDev &m, const SurfPt &sp, V3 wo)
{
	if(YG_NOT_BLEND(MT, m)) return leafAlpha<MT>(m, sp, wo);
	if(!m.is_transparent) return 1.f;
	return fminf(leafAlphaG(sc.materials[m.mat1], sp, wo), leafAlphaG(sc.materials[m.mat2], sp, wo));
}
// used by the shadow kernel on whatever transparent material the segment crosses: generic
YG_D Col matTransparency(const SceneDev &sc, const MatDev &m, const SurfPt &sp, V3 wo)
{
	if(m.type != YAFGPU_MAT_BLEND) return leafTransparency<-1>(m, sp, wo);
	const float val = m.blend_val, ival = fminf(1.f, fmaxf(0.f, 1.f - val));
	return leafTransparencyG(sc.materials[m.mat1], sp, wo) * ival + leafTransparencyG(sc.materials[m.mat2], sp, wo) * val;
}

} // namespace yg
