// K6 device code: light sampling and backgrounds.  Restates
//   PointLight::illuminate           /root/reference/src/light/light_point.cc:38-56
//   SunLight::illumSample/intersect  /root/reference/src/light/light_sun.cc:53-77
//   AreaLight::illumSample/intersect /root/reference/src/light/light_area.cc:67-160
//   MeshLight::illumSample/intersect /root/reference/src/light/light_meshlight.cc:87-222 (+ Triangle::sample triangle.cc:181-192)
//   BackgroundLight                  /root/reference/src/light/light_background.cc:122-187
//   Pdf1D::sample/dSample            /root/reference/include/utility/util_sample.h:125-163
//   ConstantBackground::eval         /root/reference/src/background/background_constant.cc:41-49
//   TextureBackground::eval          /root/reference/src/background/background_texture.cc:51-88 (+ ImageTexture::getColor, bilinear)
//   spheremap__/invSpheremap__       /root/reference/include/texture/texture.h:111-140
#pragma once
#include "materials.cuh"
#include "traverse.cuh"

namespace yg {

YG_HD bool lightIsDirac(const LightDev &L) { return L.dirac != 0; }          // Light::diracLight() (host_prep.h prepLight)
YG_HD bool lightCanIntersect(const LightDev &L) { return L.can_intersect != 0; } // Light::canIntersect()

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

// angmap__ texture.h:83-94
YG_HD void angmap(V3 p, float &u, float &v)
{
	float r = p.x * p.x + p.z * p.z;
	u = v = 0.f;
	if(r > 0.f)
	{
		const float phi_ratio = (float)(YG_M_1_PI * (double)fAcos(p.y));
		r = phi_ratio / fSqrt(r);
		u = p.x * r;
		v = p.z * r;
	}
}

// findTextureInterpolationCoordinates texture_image.cc:231-288, branch repeat = true, mirror = false: the two texel columns
// (rows) a bilinear lookup blends and the blend factor
YG_HD void texCoordsRepeat(float cf, int res, int &c1, int &c2, float &frac)
{
	c1 = ((int)cf) % res;
	if(cf > 0.f) { c2 = (c1 + 1) % res; frac = cf - (float)((int)cf); }
	else { c2 = (res - 1) % res; frac = -cf; }
}

// ImageTexture::getColor(Point3(u, v, 0)) texture_image.cc:75-87 with the factory defaults the upload insists on (clipping "repeat",
// xrepeat = yrepeat = 1, no mirror / crop / rot90 / adjustments): doMapping :139-158, then bilinearInterpolation :306-330 or
// noInterpolation :290-304 ("interpolate" = "none")
YG_HD Col texLookup(const BackgroundDev &bg, float u, float v)
{
	float tx = 0.5f * u + 0.5f, ty = 0.5f * (-v) + 0.5f;
	if(tx > 1.f) tx -= (float)((int)tx); else if(tx < 0.f) tx += (float)(1 - (int)tx);
	if(ty > 1.f) ty -= (float)((int)ty); else if(ty < 0.f) ty += (float)(1 - (int)ty);
	const int resx = bg.tex_w, resy = bg.tex_h;
	const double fx = (double)tx - floor((double)tx), fy = (double)ty - floor((double)ty);
	int x1, x2, y1, y2; float dx, dy;
	if(!bg.tex_interp)
	{
		texCoordsRepeat((float)((double)(float)resx * fx), resx, x1, x2, dx);
		texCoordsRepeat((float)((double)(float)resy * fy), resy, y1, y2, dy);
		const float *t = bg.tex + 3 * ((size_t)y1 * resx + x1);
		return mkc(t[0], t[1], t[2]);
	}
	texCoordsRepeat((float)((double)(float)resx * fx - 0.5), resx, x1, x2, dx);
	texCoordsRepeat((float)((double)(float)resy * fy - 0.5), resy, y1, y2, dy);
	const float *t11 = bg.tex + 3 * ((size_t)y1 * resx + x1), *t21 = bg.tex + 3 * ((size_t)y1 * resx + x2);
	const float *t12 = bg.tex + 3 * ((size_t)y2 * resx + x1), *t22 = bg.tex + 3 * ((size_t)y2 * resx + x2);
	const float w11 = (1 - dx) * (1 - dy), w12 = (1 - dx) * dy, w21 = dx * (1 - dy), w22 = dx * dy;
	return ((w11 * mkc(t11[0], t11[1], t11[2]) + w12 * mkc(t12[0], t12[1], t12[2])) + w21 * mkc(t21[0], t21[1], t21[2])) + w22 * mkc(t22[0], t22[1], t22[2]);
}

// Background::eval / operator(): ConstantBackground (background_constant.cc:41-49), TextureBackground (background_texture.cc:51-88)
YG_HD Col bgEval(const BackgroundDev &bg, V3 dir)
{
	if(bg.type == YAFGPU_BG_GRADIENT)
	{	// GradientBackground::eval background_gradient.cc:39-58
		Col color;
		float blend = dir.z;
		if(blend >= 0.f) color = blend * mkc(bg.szenith[0], bg.szenith[1], bg.szenith[2]) + (1.f - blend) * mkc(bg.shoriz[0], bg.shoriz[1], bg.shoriz[2]);
		else
		{
			blend = -blend;
			color = blend * mkc(bg.gzenith[0], bg.gzenith[1], bg.gzenith[2]) + (1.f - blend) * mkc(bg.ghoriz[0], bg.ghoriz[1], bg.ghoriz[2]);
		}
		if(fminf(color.r, fminf(color.g, color.b)) < 1e-6f) color = mkc(1e-5f);
		return color;
	}
	if(bg.type != YAFGPU_BG_TEXTURE) return mkc(bg.color[0], bg.color[1], bg.color[2]); // constant: col * power
	float u = 0.f, v = 0.f;
	if(bg.mapping_angular)
	{
		V3 d = dir;
		d.x = dir.x * bg.cos_r + dir.y * bg.sin_r;
		d.y = dir.x * -bg.sin_r + dir.y * bg.cos_r;
		angmap(d, u, v);
	}
	else
	{
		spheremap(dir, u, v);
		u = 2.f * u - 1.f;
		v = 2.f * v - 1.f;
		u += bg.rotation;
		if(u > 1.f) u -= 2.f;
	}
	Col ret = texLookup(bg, u, v);
	const float min_component = 1.0e-5f;
	if(ret.r < min_component) ret.r = min_component;
	if(ret.g < min_component) ret.g = min_component;
	if(ret.b < min_component) ret.b = min_component;
	return bg.power * ret;
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

// SpotLight falloff (light_spot.cc:105-114 and :139-148): false outside the cone, else the smoothstep factor (1 inside cos_start_)
YG_D bool spotFalloff(const LightDev &L, float cosa, float &v, bool &full)
{
	if(cosa < L.cos_angle) return false;
	full = cosa >= L.pdf;
	if(!full) { v = (cosa - L.cos_angle) * L.invpdf; v = v * v * (3.f - 2.f * v); }
	return true;
}

// IesLight::getAngles light_ies.cc:52-63 (u from the WORLD z of the direction, as the reference has it) and IesData::getRadiance util_ies.h:67-128.
// Table layout at L.cdf_offset: vert angles [nv], +INF, hor angles [nh], -INF, rad_map [nh][nv] (host_prep.h buildIesTables explains the -INF).
YG_D float iesRadiance(const float *__restrict__ tab, const LightDev &L, V3 dir, float costheta)
{
	float u = (dir.z >= 1.f) ? 0.f : (float)((double)fAcos(dir.z) * 57.29577951308232087684636);
	if(dir.y < 0) u = 360.f - u;
	const float v = (costheta >= 1.f) ? 0.f : (float)((double)fAcos(costheta) * 57.29577951308232087684636);
	const int nv = L.ies_nv, nh = L.ies_nh;
	const float *vert = tab + L.cdf_offset, *hor = vert + nv + 1, *rad = hor + nh + 1;
	float ang_1, ang_2;
	if(L.ies_type == 1) { ang_1 = u; ang_2 = v; }
	else
	{
		ang_1 = v; ang_2 = u;
		if(L.ies_type == 2) { ang_1 += 90; if(ang_1 > 360.f) ang_1 -= 360.f; }
	}
	if(ang_1 > 180.f && hor[nh - 1] <= 180.f) ang_1 = 360.f - ang_1;
	if(ang_1 > 90.f && hor[nh - 1] <= 90.f) ang_1 -= 90.f;
	if(ang_2 > 90.f && vert[nv - 1] <= 90.f) ang_2 -= 90.f;
	int x = 0, y = 0;
	for(int i = 0; i < nh; i++) if(hor[i] <= ang_1 && hor[i + 1] > ang_1) x = i;
	for(int i = 0; i < nv; i++) if(vert[i] <= ang_2 && vert[i + 1] > ang_2) { y = i; break; }
	float r;
	if(ang_1 == hor[x] && ang_2 == vert[y]) r = rad[x * nv + y];
	else
	{
		// a horizontal angle at or beyond the last map entry leaves x at 0 and extrapolates, as the reference does; a vertical one keeps the last cell
		const int x_2 = x + 1 < nh ? x + 1 : x, y_2 = y + 1 < nv ? y + 1 : y;
		const float d_x = x_2 == x ? 0.f : (ang_1 - hor[x]) / (hor[x_2] - hor[x]);
		const float d_y = y_2 == y ? 0.f : (ang_2 - vert[y]) / (vert[y_2] - vert[y]);
		const float rx_1 = ((1.f - d_x) * rad[x * nv + y]) + (d_x * rad[x_2 * nv + y]);
		const float rx_2 = ((1.f - d_x) * rad[x * nv + y_2]) + (d_x * rad[x_2 * nv + y_2]);
		r = ((1.f - d_y) * rx_1) + (d_y * rx_2);
	}
	return r * L.ies_max_rad;
}

// Light::illuminate of the dirac lights: PointLight light_point.cc:38-56, DirectionalLight light_directional.cc:59-81,
// SpotLight (soft_shadows off) light_spot.cc:88-120
YG_D bool lightIlluminate(const SceneDev &sc, const LightDev &L, const SurfPt &sp, Col &col, V3 &dir, float &tmax)
{
	if(L.photon_only) return false;
	if(L.type == YAFGPU_LIGHT_DIRECTIONAL)
	{
		const V3 direction = ldv(L.fnormal);
		if(L.double_sided) // finite: inside the illuminated cylinder only
		{
			const V3 vec = ldv(L.pos) - sp.p;
			const float dist = length(cross(direction, vec));
			if(dist > L.area) return false;
			tmax = dot(vec, direction);
			if(tmax <= 0.0f) return false;
		}
		else tmax = -1.0f;
		dir = direction;
		col = ldc(L.color);
		return true;
	}
	V3 ldir = ldv(L.pos) - sp.p;
	const float dist_sqr = ldir.x * ldir.x + ldir.y * ldir.y + ldir.z * ldir.z;
	const float dist = fSqrt(dist_sqr);
	if(dist == 0.0f) return false;
	const float idist_sqr = 1.f / dist_sqr;
	ldir = ldir * (1.f / dist);
	if(L.type == YAFGPU_LIGHT_SPOT)
	{
		float v = 0.f; bool full = false;
		if(!spotFalloff(L, dot(ldv(L.fnormal), ldir), v, full)) return false;
		col = full ? ldc(L.color) * idist_sqr : ldc(L.color) * (v * idist_sqr);
		tmax = dist; dir = ldir;
		return true;
	}
	if(L.type == YAFGPU_LIGHT_IES) // IesLight::illuminate light_ies.cc:65-93
	{
		const float cosa = dot(ldv(L.fnormal), ldir);
		if(cosa < L.cos_angle) return false;
		col = (ldc(L.color) * iesRadiance(sc.light_tables, L, ldir, cosa)) * idist_sqr;
		tmax = dist; dir = ldir;
		return true;
	}
	tmax = dist; dir = ldir;
	col = ldc(L.color) * idist_sqr;
	return true;
}

// sphereIntersect__ light_sphere.cc:56-68: eb, osc and the roots are formed in double (2.0, 4.0 literals)
YG_D bool sphereIntersect(V3 from, V3 dir, V3 c, float r_2, float &d_1, float &d_2)
{
	const V3 vf = from - c;
	const float ea = dot(dir, dir);
	const float eb = dot(2.0f * vf, dir);
	const float ec = dot(vf, vf) - r_2;
	float osc = (float)((double)(eb * eb) - 4.0 * (double)ea * (double)ec);
	if(osc < 0) { d_1 = fSqrt(ec / ea); return false; }
	osc = fSqrt(osc);
	d_1 = (float)((double)(-eb - osc) / (2.0 * (double)ea));
	d_2 = (float)((double)(-eb + osc) / (2.0 * (double)ea));
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
		case YAFGPU_LIGHT_BGPORTAL: // BackgroundPortalLight::illumSample light_background_portal.cc:118-150: the mesh light's sampling, single sided,
		case YAFGPU_LIGHT_MESH:     // colour = the background seen along the light direction
		{
			// sampleSurface light_meshlight.cc:87-106 / light_background_portal.cc:80-97
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
			const float amc = L.area * cos_angle;
			if(L.type == YAFGPU_LIGHT_BGPORTAL)
			{
				col = bgEval(sc.bg, ldir) * L.power;
				pdf = (float)((double)dist_sqr * YG_M_PI / (double)amc);
				return true;
			}
			col = ldc(L.color);
			pdf = (float)((double)dist_sqr * YG_M_PI / (double)((amc == 0.f) ? 1e-8f : amc));
			return true;
		}
		case YAFGPU_LIGHT_SPOT: // SpotLight::illumSample light_spot.cc:122-164
		{
			V3 ldir = ldv(L.pos) - sp.p;
			const float dist_sqr = dot(ldir, ldir);
			if(dist_sqr == 0.0f) return false;
			const float dist = fSqrt(dist_sqr);
			ldir = ldir * (1.f / dist);
			float v = 0.f; bool full = false;
			if(!spotFalloff(L, dot(ldv(L.fnormal), ldir), v, full)) return false;
			tmax = dist;
			dir = sampleCone(ldir, ldv(L.du), ldv(L.dv), L.cos_angle, s1 * L.area, s2 * L.area);
			col = full ? ldc(L.color) : ldc(L.color) * v;
			pdf = dist_sqr;
			if(pdf < 1.f) { pdf = 1.f; col = col / dist_sqr; }
			return true;
		}
		case YAFGPU_LIGHT_IES: // IesLight::illumSample light_ies.cc:95-128 (soft_shadows on): a cone of half-angle acos(cosa) around the direction to the light
		{
			V3 ldir = ldv(L.pos) - sp.p;
			const float dist_sqr = lengthSqr(ldir);
			const float dist = fSqrt(dist_sqr);
			const float idist_sqr = 1.f / dist_sqr;
			if(dist == 0.0f) return false;
			ldir = ldir * (1.f / dist);
			const float cosa = dot(ldv(L.fnormal), ldir);
			if(cosa < L.cos_angle) return false;
			tmax = dist;
			dir = sampleCone(ldir, ldv(L.du), ldv(L.dv), cosa, s1, s2);
			const float r = iesRadiance(sc.light_tables, L, dir, cosa);
			if(r == 0.f) return false;
			col = ldc(L.color) * idist_sqr;
			pdf = 1.f / r;
			return true;
		}
		case YAFGPU_LIGHT_SPHERE: // SphereLight::illumSample light_sphere.cc:70-103
		{
			V3 cdir = ldv(L.pos) - sp.p;
			const float dist_sqr = lengthSqr(cdir);
			if(dist_sqr <= L.area) return false;
			const float dist = fSqrt(dist_sqr);
			const float idist_sqr = 1.f / dist_sqr;
			const float cos_alpha = fSqrt(1.f - L.area * idist_sqr);
			cdir = cdir * (1.f / dist);
			V3 du, dv;
			createCs(cdir, du, dv);
			dir = sampleCone(cdir, du, dv, cos_alpha, s1, s2);
			float d_1, d_2;
			if(!sphereIntersect(sp.p, dir, ldv(L.pos), L.inv_area, d_1, d_2)) return false;
			tmax = d_1;
			pdf = 1.f / (2.f * (1.f - cos_alpha));
			col = ldc(L.color);
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
		case YAFGPU_LIGHT_BGPORTAL: // BackgroundPortalLight::intersect light_background_portal.cc:188-212 (clamp_intersect_ keeps its default 0)
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
			col = L.type == YAFGPU_LIGHT_BGPORTAL ? bgEval(sc.bg, dir) * L.power : ldc(L.color);
			return true;
		}
		case YAFGPU_LIGHT_SPOT: // SpotLight::intersect light_spot.cc:215-249 (soft_shadows on)
		{
			const V3 sdir = ldv(L.normal);
			const float cos_a = dot(sdir, dir);
			if(cos_a == 0.f) return false;
			t = dot(sdir, ldv(L.pos) - from) / cos_a;
			if(t < 0.f) return false;
			const V3 p = from + t * dir;
			if(dot(sdir, p - ldv(L.pos)) == 0)
			{
				if((double)dot(p, p) <= 1e-2) // sic: the hit point's squared distance from the ORIGIN
				{
					float v = 0.f; bool full = false;
					if(!spotFalloff(L, cos_a, v, full)) return false;
					col = full ? ldc(L.color) : ldc(L.color) * v;
					ipdf = 1.f / (t * t);
					return true;
				}
			}
			return false;
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
