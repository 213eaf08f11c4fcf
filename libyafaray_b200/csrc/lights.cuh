// K6 device code: light sampling and backgrounds.  Restates
//   PointLight::illuminate           /root/reference/src/light/light_point.cc:38-56
//   SunLight::illumSample/intersect  /root/reference/src/light/light_sun.cc:53-77
//   AreaLight::illumSample/intersect /root/reference/src/light/light_area.cc:67-160
//   MeshLight::illumSample/intersect /root/reference/src/light/light_meshlight.cc:87-222 (+ Triangle::sample triangle.cc:181-192)
//   BackgroundLight                  /root/reference/src/light/light_background.cc:122-187
//   Pdf1D::sample/dSample            /root/reference/include/utility/util_sample.h:125-163
//   ConstantBackground::eval         /root/reference/src/background/background_constant.cc:41-49
//   spheremap__/invSpheremap__       /root/reference/include/texture/texture.h:111-140
#pragma once
#include "materials.cuh"
#include "traverse.cuh"

namespace yg {

YG_HD bool lightIsDirac(int type) { return type == YAFGPU_LIGHT_POINT; }
YG_HD bool lightCanIntersect(int type) { return type != YAFGPU_LIGHT_POINT; } // sun, area, mesh (tree built), background

// ---- background ------------------------------------------------------------------------------------------------------
YG_HD void spheremap(V3 p, float &u, float &v)
{
	const float sqrt_r_phi = p.x * p.x + p.y * p.y;
	const float sqrt_r_theta = sqrt_r_phi + p.z * p.z;
	float phi_ratio;
	u = 0.f; v = 0.f;
	if(sqrt_r_phi > 0.f)
	{
		if(p.y < 0.f) phi_ratio = (float)((YG_M_2PI - (double)fAcos(p.x / fSqrt(sqrt_r_phi))) * 0.15915494309189533577);
		else phi_ratio = (float)((double)fAcos(p.x / fSqrt(sqrt_r_phi)) * 0.15915494309189533577);
		u = 1.f - phi_ratio;
	}
	v = (float)(1.0 - ((double)fAcos(p.z / fSqrt(sqrt_r_theta)) * YG_M_1_PI)); // 1.f - (float * double): evaluated in double
}
YG_HD V3 invSpheremap(float u, float v)
{
	const float theta = mulD(v, YG_M_PI);
	const float phi = -mulD(u, YG_M_2PI);
	const float costheta = fCos(theta), sintheta = fSin(theta);
	const float cosphi = fCos(phi), sinphi = fSin(phi);
	return mk(sintheta * cosphi, sintheta * sinphi, -costheta);
}

// Background::eval / operator() for the supported types
YG_HD Col bgEval(const BackgroundDev &bg, V3 dir)
{
	(void)dir;
	return mkc(bg.color[0], bg.color[1], bg.color[2]); // constant: col * power
}

// ---- Pdf1D -------------------------------------------------------------------------------------------------------------
// std::lower_bound(cdf, cdf + count + 1, u) - cdf
YG_HD int lowerBound(const float *cdf, int n, float u)
{
	int lo = 0, len = n;
	while(len > 0)
	{
		const int half = len >> 1;
		if(cdf[lo + half] < u) { lo += half + 1; len -= half + 1; }
		else len = half;
	}
	return lo;
}
YG_HD float pdf1dSample(const float *func, const float *cdf, int count, float inv_integral, float u, float &pdf)
{
	int index = lowerBound(cdf, count + 1, u) - 1;
	if(index < 0) index = 0;
	const float delta = (u - cdf[index]) / (cdf[index + 1] - cdf[index]);
	pdf = func[index] * inv_integral;
	return (float)index + delta;
}
YG_HD int pdf1dDSample(const float *cdf, int count, float u)
{
	if(u == 0.f) return 0;
	int index = lowerBound(cdf, count + 1, u) - 1;
	if(index < 0) index = 0;
	return index;
}

// ---- background light tables (layout written by host_prep.h buildBgLightTables) ----------------------------------------
// at tab[L.cdf_offset]: v_func[nv], v_cdf[nv+1], then per row r: {bits(func_off), bits(count), integral, inv_integral} x nv,
// then the row data (func[count], cdf[count+1]) at func_off (absolute offsets into the table array).
struct BgRow { int func_off, count; float integral, inv_integral; };
YG_HD BgRow bgRow(const float *tab, const LightDev &L, int iv)
{
	const float *r = tab + L.cdf_offset + (2 * L.cdf_count + 1) + 4 * iv;
	BgRow b; b.func_off = f2i(r[0]); b.count = f2i(r[1]); b.integral = r[2]; b.inv_integral = r[3];
	return b;
}
YG_HD float bgSinSample(float s) { return fSin(mulD(s, YG_M_PI)); }
YG_HD float bgClampZero(float v) { return v > 0.f ? 1.f / v : 0.f; }
YG_HD float bgCalcPdf(float p0, float p1, float s, bool inv)
{
	if(inv) return fmaxf(0.000001f, (float)YG_M_2PI * bgSinSample(s) * bgClampZero(p0 * p1));
	return fmaxf(0.000001f, (p0 * p1) * (float)0.15915494309189533577 * bgClampZero(bgSinSample(s)));
}
YG_HD int bgClampSample(float s, int m) { int i = (int)s; return i < 0 ? 0 : (i > m - 1 ? m - 1 : i); }
YG_HD float bgCalcFromSample(const float *tab, const LightDev &L, float s1, float s2, float &u, float &v, bool inv)
{
	const int nv = L.cdf_count;
	const float *v_func = tab + L.cdf_offset, *v_cdf = v_func + nv;
	float pdf_1 = 0.f, pdf_2 = 0.f;
	v = pdf1dSample(v_func, v_cdf, nv, 1.f / L.cdf_integral, s2, pdf_2);
	const int iv = bgClampSample(v + 0.4999f, nv);
	const BgRow row = bgRow(tab, L, iv);
	u = pdf1dSample(tab + row.func_off, tab + row.func_off + row.count, row.count, row.inv_integral, s1, pdf_1);
	u *= 1.f / (float)row.count;
	v *= 1.f / (float)nv;
	return bgCalcPdf(pdf_1, pdf_2, v, inv);
}
YG_HD float bgCalcFromDir(const float *tab, const LightDev &L, V3 dir, float &u, float &v, bool inv)
{
	const int nv = L.cdf_count;
	const float *v_func = tab + L.cdf_offset;
	spheremap(dir, u, v);
	const int iv = bgClampSample(v * (float)nv + 0.4999f, nv);
	const BgRow row = bgRow(tab, L, iv);
	const int iu = bgClampSample(u * (float)row.count + 0.4999f, row.count);
	const float pdf_1 = tab[row.func_off + iu] * row.inv_integral;
	const float pdf_2 = v_func[iv] * (1.f / L.cdf_integral);
	return bgCalcPdf(pdf_1, pdf_2, v, inv);
}

// ---- lights ------------------------------------------------------------------------------------------------------------

// PointLight::illuminate light_point.cc:38-56
YG_D bool lightIlluminate(const LightDev &L, const SurfPt &sp, Col &col, V3 &dir, float &tmax)
{
	if(L.photon_only) return false;
	V3 ldir = ldv(L.pos) - sp.p;
	const float dist_sqr = ldir.x * ldir.x + ldir.y * ldir.y + ldir.z * ldir.z;
	const float dist = fSqrt(dist_sqr);
	if(dist == 0.0f) return false;
	const float idist_sqr = 1.f / dist_sqr;
	ldir = ldir * (1.f / dist);
	tmax = dist; dir = ldir;
	col = ldc(L.color) * idist_sqr;
	return true;
}

// Light::illumSample for sun / area / mesh / background
YG_D bool lightIllumSample(const SceneDev &sc, const LightDev &L, const SurfPt &sp, float s1, float s2, Col &col, float &pdf, V3 &dir, float &tmax)
{
	if(L.photon_only) return false;
	switch(L.type)
	{
		case YAFGPU_LIGHT_SUN:
			dir = sampleCone(ldv(L.pos), ldv(L.du), ldv(L.dv), L.cos_angle, s1, s2);
			tmax = -1.f;
			col = ldc(L.col_pdf);
			pdf = L.pdf;
			return true;
		case YAFGPU_LIGHT_AREA:
		{
			const V3 p = (ldv(L.pos) + s1 * ldv(L.to_x)) + s2 * ldv(L.to_y);
			V3 ldir = p - sp.p;
			const float dist_sqr = lengthSqr(ldir);
			const float dist = fSqrt(dist_sqr);
			if(dist <= 0.0f) return false;
			ldir = ldir * (1.f / dist);
			const float cos_angle = dot(ldir, ldv(L.fnormal));
			if(cos_angle <= 0) return false;
			tmax = dist; dir = ldir;
			col = ldc(L.color);
			pdf = (float)((double)dist_sqr * YG_M_PI / (double)(L.area * cos_angle));
			return true;
		}
		case YAFGPU_LIGHT_MESH:
		{
			// sampleSurface light_meshlight.cc:87-106
			const float *cdf = sc.light_tables + L.cdf_offset; // cdf[count+1]
			const int prim = pdf1dDSample(cdf, L.cdf_count, s1);
			if(prim >= L.cdf_count) return false;
			float ss_1, delta = cdf[prim + 1];
			if(prim > 0) { delta -= cdf[prim]; ss_1 = (s1 - cdf[prim]) / delta; }
			else ss_1 = s1 / delta;
			// Triangle::sample triangle.cc:181-192
			const TriShade *ts = sc.tris + (L.tri_first + prim);
			const V3 a = ldv(sc.points + 3 * ts->v[0]), b = ldv(sc.points + 3 * ts->v[1]), c = ldv(sc.points + 3 * ts->v[2]);
			const float su_1 = fSqrt(ss_1);
			const float u = 1.f - su_1, v = s2 * su_1;
			const V3 p = (u * a + v * b) + ((1.f - u) - v) * c;
			const V3 n = ldv(ts->ng);
			V3 ldir = p - sp.p;
			const float dist_sqr = lengthSqr(ldir);
			const float dist = fSqrt(dist_sqr);
			if(dist <= 0.0f) return false;
			ldir = ldir * (1.f / dist);
			float cos_angle = -dot(ldir, n);
			if(cos_angle <= 0)
			{
				if(L.double_sided) cos_angle = -cos_angle;
				else return false;
			}
			tmax = dist; dir = ldir;
			col = ldc(L.color);
			const float amc = L.area * cos_angle;
			pdf = (float)((double)dist_sqr * YG_M_PI / (double)((amc == 0.f) ? 1e-8f : amc));
			return true;
		}
		case YAFGPU_LIGHT_BACKGROUND:
		{
			float u = 0.f, v = 0.f;
			tmax = -1.0f;
			pdf = bgCalcFromSample(sc.light_tables, L, s1, s2, u, v, false);
			dir = invSpheremap(u, v);
			col = bgEval(sc.bg, dir);
			return true;
		}
		default: return false;
	}
}

// triIntersect__ light_area.cc:118-137
YG_D bool areaTriIntersect(V3 a, V3 b, V3 c, V3 from, V3 dir, float &t)
{
	const V3 edge_1 = b - a, edge_2 = c - a;
	const V3 pvec = cross(dir, edge_2);
	const float det = dot(edge_1, pvec);
	if(det == 0.0f) return false;
	const float inv_det = (float)(1.0 / (double)det);
	const V3 tvec = from - a;
	const float u = dot(tvec, pvec) * inv_det;
	if(u < 0.0f || u > 1.0f) return false;
	const V3 qvec = cross(tvec, edge_1);
	const float v = dot(dir, qvec) * inv_det;
	if((v < 0.0f) || ((u + v) > 1.0f)) return false;
	t = dot(edge_2, qvec) * inv_det;
	return true;
}

// Light::intersect(ray, t, col, ipdf); ray.tmax is -1 (default-constructed b_ray, integrator_montecarlo.cc:283)
YG_D bool lightIntersect(const SceneDev &sc, const LightDev &L, V3 from, V3 dir, float tmin, float &t, Col &col, float &ipdf)
{
	switch(L.type)
	{
		case YAFGPU_LIGHT_SUN:
		{
			const float cosine = dot(dir, ldv(L.pos));
			if(cosine < L.cos_angle) return false;
			col = ldc(L.col_pdf);
			t = -1;
			ipdf = L.invpdf;
			return true;
		}
		case YAFGPU_LIGHT_AREA:
		{
			const float cos_angle = dot(dir, ldv(L.fnormal));
			if(cos_angle <= 0) return false;
			if(!areaTriIntersect(ldv(L.pos), ldv(L.c2), ldv(L.c3), from, dir, t))
			{
				if(!areaTriIntersect(ldv(L.pos), ldv(L.c3), ldv(L.c4), from, dir, t)) return false;
			}
			if(!(t > 1.0e-10f)) return false;
			col = ldc(L.color);
			ipdf = (float)((double)(1.f / (t * t) * L.area * cos_angle) * YG_M_1_PI);
			return true;
		}
		case YAFGPU_LIGHT_MESH:
		{
			if(L.accel < 0) return false;
			RayD r; r.o = from; r.d = dir; r.tmin = tmin; r.tmax = -1.f;
			const HitD h = traceClosest<false>(sc.light_accels[L.accel], r, nullptr);
			if(h.tri < 0) return false;
			t = h.t;
			const V3 n = ldv(sc.tris[L.tri_first + h.tri].ng);
			float cos_angle = dot(dir, -n);
			if(cos_angle <= 0)
			{
				if(L.double_sided) cos_angle = fabsf(cos_angle);
				else return false;
			}
			const float idist_sqr = 1.f / (t * t);
			ipdf = idist_sqr * L.area * cos_angle * (float)(1.f / YG_M_PI);
			col = ldc(L.color);
			return true;
		}
		case YAFGPU_LIGHT_BACKGROUND:
		{
			float u = 0.f, v = 0.f;
			V3 abs_dir = dir;
			if(L.abs_intersect) abs_dir = -abs_dir;
			ipdf = bgCalcFromDir(sc.light_tables, L, abs_dir, u, v, true);
			const V3 d2 = invSpheremap(u, v);
			col = bgEval(sc.bg, d2);
			col = clampProportional(col, L.clamp_intersect);
			t = -1.f; // b_ray.tmax_ stays at its default (infinite)
			return true;
		}
		default: return false;
	}
}

} // namespace yg
