// Shader-node graphs and image textures on the device.
//
// Restates, operation for operation (float32, no FMA contraction, double where the reference promotes):
//   Triangle::getSurface (orco / uv / dPdU part)   /root/reference/src/common/triangle.cc:47-131
//   TextureMapperNode                               /root/reference/src/shader/shader_node_basic.cc:28-342
//   ValueNode, MixNode and its eight subclasses     /root/reference/src/shader/shader_node_basic.cc:426-705
//   LayerNode, textureRgbBlend__/ValueBlend__       /root/reference/src/shader/shader_node_layer.cc:31-148, include/shader/shader_node.h:115-223
//   ImageTexture::getColor/doMapping/interpolation  /root/reference/src/texture/texture_image.cc:75-375
//   Texture::applyAdjustments, getFloat             /root/reference/include/texture/texture.h:49-52,232-275
//   Rgb colour-space / HSV helpers                  /root/reference/include/common/color.h:352-410,467-507
//   NodeMaterial::evalNodes/evalBump, applyBump     /root/reference/src/material/material_node.cc:79-139, src/common/material.cc:77-84
//
// The reference evaluates a material's nodes once per hit in initBsdf() into a per-thread "node stack" and its BSDF functions then
// read colours/scalars from that stack instead of the material constants.  None of the implemented nodes is view dependent, so here the
// stack lives only inside evalMaterialNodes(): it resolves every bound "<slot>_shader" into a thread-local copy of the material record,
// and the unchanged BSDF code runs on that copy.
#pragma once
#include "materials.cuh"

namespace yg {

struct Rgba4 { float r, g, b, a; };
YG_D Rgba4 mk4(float r, float g, float b, float a) { Rgba4 c; c.r = r; c.g = g; c.b = b; c.a = a; return c; }
YG_D Rgba4 operator*(float f, Rgba4 c) { return mk4(f * c.r, f * c.g, f * c.b, f * c.a); }              // color.h:300-308
YG_D Rgba4 operator*(Rgba4 c, float f) { return mk4(f * c.r, f * c.g, f * c.b, f * c.a); }
YG_D Rgba4 operator*(Rgba4 a, Rgba4 b) { return mk4(a.r * b.r, a.g * b.g, a.b * b.b, a.a * b.a); }
YG_D Rgba4 operator+(Rgba4 a, Rgba4 b) { return mk4(a.r + b.r, a.g + b.g, a.b + b.b, a.a + b.a); }
YG_D Rgba4 operator-(Rgba4 a, Rgba4 b) { return mk4(a.r - b.r, a.g - b.g, a.b - b.b, a.a - b.a); }
YG_D float col2Bri(Rgba4 c) { return (0.2126f * c.r + 0.7152f * c.g) + 0.0722f * c.b; }                  // color.h:60
YG_D Col rgbOf(Rgba4 c) { return mkc(c.r, c.g, c.b); }

struct NodeRes { Rgba4 col; float f; };

// extra surface-point fields only shader nodes read
struct SurfTex
{
	float u, v; int has_uv;
	V3 orco_p, orco_ng;
	V3 ds_du, ds_dv;
};

// Triangle::getSurface triangle.cc:47-131 — the part getSurface() in materials.cuh leaves out.  Call after getSurface (needs sp.n/nu/nv).
YG_D void getSurfaceTex(const SceneDev &sc, int tri, float b1, float b2, const SurfPt &sp, SurfTex &sx)
{
	const TriShade *ts = sc.tris + tri;
	const ObjDev ob = sc.objects[ts->obj];
	const float u = (1 - b1) - b2, v = b1, w = b2;
	const V3 p0 = ldv(sc.points + 3 * (size_t)ts->v[0]), p1 = ldv(sc.points + 3 * (size_t)ts->v[1]), p2 = ldv(sc.points + 3 * (size_t)ts->v[2]);
	if(ob.has_orco)
	{
		const V3 o0 = ldv(sc.points + 3 * (size_t)(ts->v[0] + 1)), o1 = ldv(sc.points + 3 * (size_t)(ts->v[1] + 1)), o2 = ldv(sc.points + 3 * (size_t)(ts->v[2] + 1));
		sx.orco_p = (u * o0 + v * o1) + w * o2;
		sx.orco_ng = normalize(cross(o1 - o0, o2 - o0));
	}
	else { sx.orco_p = sp.p; sx.orco_ng = sp.ng; }
	V3 dp_du, dp_dv;
	if(ob.has_uv)
	{
		const float *t1 = sc.uv + 2 * (size_t)ts->uv[0], *t2 = sc.uv + 2 * (size_t)ts->uv[1], *t3 = sc.uv + 2 * (size_t)ts->uv[2];
		const float u1 = t1[0], v1 = t1[1], u2 = t2[0], v2 = t2[1], u3 = t3[0], v3 = t3[1];
		sx.u = (u * u1 + v * u2) + w * u3;
		sx.v = (u * v1 + v * v2) + w * v3;
		const float du_1 = u1 - u3, du_2 = u2 - u3, dv_1 = v1 - v3, dv_2 = v2 - v3;
		const float det = du_1 * dv_2 - dv_1 * du_2;
		if(fabsf(det) > 1e-30f)
		{
			const float invdet = 1.f / det;
			const V3 dp_1 = p0 - p2, dp_2 = p1 - p2;
			dp_du = (dv_2 * dp_1 - dv_1 * dp_2) * invdet;
			dp_dv = (du_1 * dp_2 - du_2 * dp_1) * invdet;
		}
		else { dp_du = p1 - p0; dp_dv = p2 - p1; }
	}
	else { dp_du = p1 - p0; dp_dv = p2 - p1; sx.u = 0.f; sx.v = 0.f; }
	dp_du = normalize(dp_du); dp_dv = normalize(dp_dv);
	sx.ds_du = mk(dot(sp.nu, dp_du), dot(sp.nv, dp_du), dot(sp.n, dp_du));
	sx.ds_dv = mk(dot(sp.nu, dp_dv), dot(sp.nv, dp_dv), dot(sp.n, dp_dv));
	sx.has_uv = ob.has_uv;
}

// ---- colour helpers -----------------------------------------------------------------------------------------------------
YG_D float sRgbFromLinear(float v) { return v <= 0.0031308f ? v * 12.92f : (1.055f * fPow(v, 0.416667f)) - 0.055f; } // color.h:359-364
// Rgb::colorSpaceFromLinearRgb color.h:388-410 (alpha untouched)
YG_D Rgba4 colorSpaceFromLinear(Rgba4 c, int cs, float gamma)
{
	if(cs == YAFGPU_CS_SRGB) { c.r = sRgbFromLinear(c.r); c.g = sRgbFromLinear(c.g); c.b = sRgbFromLinear(c.b); }
	else if(cs == YAFGPU_CS_XYZ_D65)
	{	// xyz_d_65_from_linear_rgb__ color.h:345-350
		const float r = c.r, g = c.g, b = c.b;
		c.r = (0.412400f * r + 0.357600f * g) + 0.180500f * b;
		c.g = (0.212600f * r + 0.715200f * g) + 0.072200f * b;
		c.b = (0.019300f * r + 0.119200f * g) + 0.950500f * b;
	}
	else if(cs == YAFGPU_CS_RAW_MANUAL_GAMMA && gamma != 1.f)
	{
		if(gamma <= 0.f) gamma = 1.0e-2f;
		const float ig = 1.f / gamma;
		c.r = fPow(c.r, ig); c.g = fPow(c.g, ig); c.b = fPow(c.b, ig);
	}
	return c;
}
YG_D void rgbToHsv(Rgba4 c, float &h, float &s, float &v)
{	// color.h:467-487
	const float r1 = fmaxf(c.r, 0.f), g1 = fmaxf(c.g, 0.f), b1 = fmaxf(c.b, 0.f);
	const float mx = fmaxf(fmaxf(r1, g1), b1), mn = fminf(fminf(r1, g1), b1);
	const float range = mx - mn;
	v = mx;
	if(fabsf(range) < 1.0e-6f) { h = 0.f; s = 0.f; }
	else if(mx == r1) { h = fmodf((g1 - b1) / range, 6.f); s = range / fmaxf(v, 1.0e-6f); }
	else if(mx == g1) { h = ((b1 - r1) / range) + 2.f; s = range / fmaxf(v, 1.0e-6f); }
	else if(mx == b1) { h = ((r1 - g1) / range) + 4.f; s = range / fmaxf(v, 1.0e-6f); }
	else { h = 0.f; s = 0.f; v = 0.f; }
	if(h < 0.f) h += 6.f;
}
YG_D void hsvToRgb(Rgba4 &c, float h, float s, float v)
{	// color.h:489-507
	const float ch = v * s;
	const float x = ch * (1.f - fabsf(fmodf(h, 2.f) - 1.f));
	const float m = v - ch;
	float r1 = 0.f, g1 = 0.f, b1 = 0.f;
	if(h >= 0.f && h < 1.f) { r1 = ch; g1 = x; }
	else if(h >= 1.f && h < 2.f) { r1 = x; g1 = ch; }
	else if(h >= 2.f && h < 3.f) { g1 = ch; b1 = x; }
	else if(h >= 3.f && h < 4.f) { g1 = x; b1 = ch; }
	else if(h >= 4.f && h < 5.f) { r1 = x; b1 = ch; }
	else if(h >= 5.f && h < 6.f) { r1 = ch; b1 = x; }
	c.r = r1 + m; c.g = g1 + m; c.b = b1 + m;
}
YG_D void clampRgb0(Rgba4 &c) { if(c.r < 0.f) c.r = 0.f; if(c.g < 0.f) c.g = 0.f; if(c.b < 0.f) c.b = 0.f; }

// Texture::applyColorAdjustments texture.h:226-252: channel factors, saturation and hue (the procedural textures apply intensity / contrast to their scalar instead)
YG_D Rgba4 texAdjustColor(const TexDev &t, Rgba4 c);
// Texture::applyAdjustments texture.h:203-258 = applyColorAdjustments(applyIntensityContrastAdjustments(c))
YG_D Rgba4 texAdjust(const TexDev &t, Rgba4 c)
{
	if(!t.adj_set) return c;
	Rgba4 ret = c;
	if(t.adj_intensity != 1.f || t.adj_contrast != 1.f)
	{
		ret.r = (c.r - 0.5f) * t.adj_contrast + t.adj_intensity - 0.5f;
		ret.g = (c.g - 0.5f) * t.adj_contrast + t.adj_intensity - 0.5f;
		ret.b = (c.b - 0.5f) * t.adj_contrast + t.adj_intensity - 0.5f;
	}
	if(t.adj_clamp) clampRgb0(ret);
	return texAdjustColor(t, ret);
}
YG_D Rgba4 texAdjustColor(const TexDev &t, Rgba4 c)
{
	if(!t.adj_set) return c;
	Rgba4 ret = c;
	if(t.adj_mult[0] != 1.f) ret.r *= t.adj_mult[0];
	if(t.adj_mult[1] != 1.f) ret.g *= t.adj_mult[1];
	if(t.adj_mult[2] != 1.f) ret.b *= t.adj_mult[2];
	if(t.adj_clamp) clampRgb0(ret);
	if(t.adj_saturation != 1.f || t.adj_hue != 0.f)
	{
		float h = 0.f, s = 0.f, v = 0.f;
		rgbToHsv(ret, h, s, v);
		s *= t.adj_saturation;
		h += t.adj_hue;
		if(h < 0.f) h += 6.f; else if(h > 6.f) h -= 6.f;
		hsvToRgb(ret, h, s, v);
		if(t.adj_clamp) clampRgb0(ret);
	}
	return ret;
}
// Texture::applyIntensityContrastAdjustments(float) texture.h:260-273
YG_D float texAdjustFloat(const TexDev &t, float f)
{
	if(!t.adj_set) return f;
	float ret = f;
	if(t.adj_intensity != 1.f || t.adj_contrast != 1.f) ret = (f - 0.5f) * t.adj_contrast + t.adj_intensity - 0.5f;
	if(t.adj_clamp) { if(ret < 0.f) ret = 0.f; else if(ret > 1.f) ret = 1.f; }
	return ret;
}

// ---- ImageTexture ---------------------------------------------------------------------------------------------------------
// ImageTexture::doMapping texture_image.cc:139-222; returns true when the point is clipped away
YG_D bool imgDoMapping(const TexDev &t, float &x, float &y, float &z)
{
	bool outside = false;
	x = 0.5f * x + 0.5f; y = 0.5f * y + 0.5f; z = 0.5f * z + 0.5f;
	if(t.clip == YAFGPU_TEXCLIP_REPEAT)
	{
		if(t.xrepeat > 1) x *= (float)t.xrepeat;
		if(t.yrepeat > 1) y *= (float)t.yrepeat;
		if(t.mirror_x && ((int)ceilf(x)) % 2 == 0) x = -x;
		if(t.mirror_y && ((int)ceilf(y)) % 2 == 0) y = -y;
		if(x > 1.f) x -= (float)((int)x); else if(x < 0.f) x += (float)(1 - (int)x);
		if(y > 1.f) y -= (float)((int)y); else if(y < 0.f) y += (float)(1 - (int)y);
	}
	if(t.cropx) x = t.cropminx + x * (t.cropmaxx - t.cropminx);
	if(t.cropy) y = t.cropminy + y * (t.cropmaxy - t.cropminy);
	if(t.rot90) { const float s = x; x = y; y = s; }
	switch(t.clip)
	{
		case YAFGPU_TEXCLIP_CLIPCUBE:
			if((x < 0) || (x > 1) || (y < 0) || (y > 1) || (z < -1) || (z > 1)) outside = true;
			break;
		case YAFGPU_TEXCLIP_CHECKER:
		{
			const int xs = (int)floor((double)x), ys = (int)floor((double)y);
			x -= (float)xs; y -= (float)ys;
			if(!t.checker_odd && !((xs + ys) & 1)) { outside = true; break; }
			if(!t.checker_even && ((xs + ys) & 1)) { outside = true; break; }
			if((double)t.checker_dist < 1.0)
			{
				x = (float)(((double)x - 0.5) / (1.0 - (double)t.checker_dist) + 0.5);
				y = (float)(((double)y - 0.5) / (1.0 - (double)t.checker_dist) + 0.5);
			}
		}	// falls through to clip
		case YAFGPU_TEXCLIP_CLIP:
			if((x < 0) || (x > 1) || (y < 0) || (y > 1)) outside = true;
			break;
		case YAFGPU_TEXCLIP_EXTEND:
			if(x > 0.99999f) x = 0.99999f; else if(x < 0) x = 0;
			if(y > 0.99999f) y = 0.99999f; else if(y < 0) y = 0;
		default: outside = false;
	}
	return outside;
}

// ImageTexture::findTextureInterpolationCoordinates texture_image.cc:231-288
YG_D void texInterpCoords(int &c0, int &c1, int &c2, int &c3, float &frac, float cf, int res, bool repeat, bool mirror)
{
	if(repeat)
	{
		c1 = ((int)cf) % res;
		if(mirror)
		{
			if(cf < 0.f) { c0 = 1 % res; c2 = c1; c3 = c0; frac = -cf; }
			else if(cf >= (float)res - 1.f) { c0 = (res + res - 1) % res; c2 = c1; c3 = c0; frac = cf - (float)((int)cf); }
			else
			{
				c0 = (res + c1 - 1) % res;
				c2 = c1 + 1; if(c2 >= res) c2 = (res + res - c2) % res;
				c3 = c1 + 2; if(c3 >= res) c3 = (res + res - c3) % res;
				frac = cf - (float)((int)cf);
			}
		}
		else
		{
			if(cf > 0.f) { c0 = (res + c1 - 1) % res; c2 = (c1 + 1) % res; c3 = (c1 + 2) % res; frac = cf - (float)((int)cf); }
			else { c0 = 1 % res; c2 = (res - 1) % res; c3 = (res - 2) % res; frac = -cf; }
		}
	}
	else
	{
		c1 = max(0, min(res - 1, (int)cf));
		if(cf > 0.f) c2 = min(res - 1, c1 + 1); else c2 = 0;
		c0 = max(0, c1 - 1);
		c3 = min(res - 1, c2 + 1);
		frac = (float)((double)cf - floor((double)cf));
	}
}

YG_D Rgba4 texel(const TexDev &t, int x, int y)
{
	const float4 q = __ldg(t.tex + ((size_t)y * t.w + x));
	return mk4(q.x, q.y, q.z, q.w);
}
// cubicInterpolate__<Rgba> util_interpolation.h:48-63 (mu is a double parameter; Rgba arithmetic narrows it to float)
YG_D Rgba4 cubicInterp(Rgba4 y0, Rgba4 y1, Rgba4 y2, Rgba4 y3, float dx)
{
	const float mu = dx, mu2 = (float)((double)dx * (double)dx);
	const Rgba4 a0 = ((y3 - y2) - y0) + y1;
	const Rgba4 a1 = (y0 - y1) - a0;
	const Rgba4 a2 = y2 - y0;
	const Rgba4 m2 = mk4(mu2, mu2, mu2, mu2);
	return (((a0 * mu) * m2 + a1 * m2) + a2 * mu) + y1;
}

// ImageTexture::interpolateImage :49-73 without mipmaps (noInterpolation :290-304, bilinear :306-330, bicubic :332-375)
YG_D Rgba4 imgInterpolate(const TexDev &t, float px, float py)
{
	const int resx = t.w, resy = t.h;
	const bool repeat = t.clip == YAFGPU_TEXCLIP_REPEAT;
	const double fx = (double)px - floor((double)px), fy = (double)py - floor((double)py);
	int x0, x1, x2, x3, y0, y1, y2, y3; float dx, dy;
	if(t.interp == YAFGPU_TEXINTERP_NONE)
	{
		texInterpCoords(x0, x1, x2, x3, dx, (float)((double)(float)resx * fx), resx, repeat, t.mirror_x);
		texInterpCoords(y0, y1, y2, y3, dy, (float)((double)(float)resy * fy), resy, repeat, t.mirror_y);
		return texel(t, x1, y1);
	}
	texInterpCoords(x0, x1, x2, x3, dx, (float)((double)(float)resx * fx - 0.5), resx, repeat, t.mirror_x);
	texInterpCoords(y0, y1, y2, y3, dy, (float)((double)(float)resy * fy - 0.5), resy, repeat, t.mirror_y);
	if(t.interp == YAFGPU_TEXINTERP_BICUBIC)
	{
		const Rgba4 cy0 = cubicInterp(texel(t, x0, y0), texel(t, x1, y0), texel(t, x2, y0), texel(t, x3, y0), dx);
		const Rgba4 cy1 = cubicInterp(texel(t, x0, y1), texel(t, x1, y1), texel(t, x2, y1), texel(t, x3, y1), dx);
		const Rgba4 cy2 = cubicInterp(texel(t, x0, y2), texel(t, x1, y2), texel(t, x2, y2), texel(t, x3, y2), dx);
		const Rgba4 cy3 = cubicInterp(texel(t, x0, y3), texel(t, x1, y3), texel(t, x2, y3), texel(t, x3, y3), dx);
		return cubicInterp(cy0, cy1, cy2, cy3, dy);
	}
	const Rgba4 c11 = texel(t, x1, y1), c21 = texel(t, x2, y1), c12 = texel(t, x1, y2), c22 = texel(t, x2, y2);
	const float w11 = (1 - dx) * (1 - dy), w12 = (1 - dx) * dy, w21 = dx * (1 - dy), w22 = dx * dy;
	return ((w11 * c11 + w12 * c12) + w21 * c21) + w22 * c22;
}

// ---- ColorRamp::getColorInterpolated color_ramp.cc:84-164 --------------------------------------------------------------------------------------
YG_D void rgbToHsl(Rgba4 c, float &h, float &s, float &l)
{	// color.h:509-529
	const float r1 = fmaxf(c.r, 0.f), g1 = fmaxf(c.g, 0.f), b1 = fmaxf(c.b, 0.f);
	const float mx = fmaxf(fmaxf(r1, g1), b1), mn = fminf(fminf(r1, g1), b1);
	const float range = mx - mn;
	l = 0.5f * (mx + mn);
	if(fabsf(range) < 1.0e-6f) { h = 0.f; s = 0.f; }
	else if(mx == r1) { h = fmodf((g1 - b1) / range, 6.f); s = range / fmaxf((1.f - fabsf((2.f * l) - 1)), 1.0e-6f); }
	else if(mx == g1) { h = ((b1 - r1) / range) + 2.f; s = range / fmaxf((1.f - fabsf((2.f * l) - 1)), 1.0e-6f); }
	else if(mx == b1) { h = ((r1 - g1) / range) + 4.f; s = range / fmaxf((1.f - fabsf((2.f * l) - 1)), 1.0e-6f); }
	else { h = 0.f; s = 0.f; l = 0.f; }
	if(h < 0.f) h += 6.f;
}
// interpolationLinear__ color_ramp.cc:58-82: item 1 is the lower bound (the item at or behind `pos`), item 2 the one in front of it
YG_D float rampLerp(float pos, float v1, float p1, float v2, float p2)
{
	if(pos == p1 || p1 == p2) return v1;
	if(pos == p2) return v2;
	return v1 + (((pos - p1) / (p2 - p1)) * (v2 - v1));
}
YG_D Rgba4 rampColor(const TexDev &t, float pos)
{
	const float *rp = t.ramp;
	auto item = [&](int k) { return mk4(rp[5 * k], rp[5 * k + 1], rp[5 * k + 2], rp[5 * k + 3]); };
	if(pos < 0.f) return item(0);
	if(pos > 1.f) return item(t.n_ramp - 1);
	int cur = 0;
	while(cur < t.n_ramp - 1 && rp[5 * cur + 4] < pos) ++cur;   // std::lower_bound over the positions (prepTexture guarantees a last position >= 1)
	// pos sits exactly on the first item (position <= 0 <= pos): every interpolation below returns item 1's value, whatever std::prev(begin) holds in the reference
	const int prev = cur > 0 ? cur - 1 : cur;
	const Rgba4 c1 = item(cur), c2 = item(prev);
	const float p1 = rp[5 * cur + 4], p2 = rp[5 * prev + 4];
	if(t.ramp_mode == YAFGPU_RAMP_RGB)
	{
		if(!t.ramp_interp) return c1;
		if(pos == p1 || p1 == p2) return c1;
		if(pos == p2) return c2;
		const float f = (pos - p1) / (p2 - p1);
		return c1 + (f * (c2 - c1));
	}
	float h1 = 0.f, s1 = 0.f, v1 = 0.f, h2 = 0.f, s2 = 0.f, v2 = 0.f;
	if(t.ramp_mode == YAFGPU_RAMP_HSV) { rgbToHsv(c1, h1, s1, v1); rgbToHsv(c2, h2, s2, v2); }
	else { rgbToHsl(c1, h1, s1, v1); rgbToHsl(c2, h2, s2, v2); }
	const float s = rampLerp(pos, s1, p1, s2, p2), v = rampLerp(pos, v1, p1, v2, p2), a = rampLerp(pos, c1.a, p1, c2.a, p2);
	const int hi = t.ramp_hue;
	if(hi == YAFGPU_RAMPHUE_CW && h1 < h2) h1 += 6.f;
	else if(hi == YAFGPU_RAMPHUE_CCW && h1 > h2) h2 += 6.f;
	else if(hi == YAFGPU_RAMPHUE_NEAR && h1 < h2 && (h2 - h1) > 3.f) h1 += 6.f;
	else if(hi == YAFGPU_RAMPHUE_NEAR && h1 > h2 && (h2 - h1) < -3.f) h2 += 6.f;
	else if(hi == YAFGPU_RAMPHUE_FAR && h1 < h2 && (h2 - h1) < 3.f) h1 += 6.f;
	else if(hi == YAFGPU_RAMPHUE_FAR && h1 > h2 && (h2 - h1) > -3.f) h2 += 6.f;
	float h = rampLerp(pos, h1, p1, h2, p2);
	if(h < 0.f) h += 6.f; else if(h > 6.f) h -= 6.f;
	Rgba4 result = mk4(0.f, 0.f, 0.f, 1.f);
	hsvToRgb(result, h, s, v);   // the HSL mode converts back with hsvToRgb as well (color_ramp.cc:158)
	result.a = a;
	return result;
}

// ---- noise bases and the procedural textures over them ------------------------------------------------------------------------
// NoiseGenerator::operator() of the five families the factories can construct (noise_generator.cc:182-630), evaluated with the reference's own
// lattice tables (NoiseTables).  Every double-typed literal of the reference is kept as one: `0.5 + 0.5 * nv`, `1.0 - 3.0 * c + 2.0 * c * o`,
// `t * t * (3.0 - 2.0 * t)` are double expressions narrowed on assignment.
YG_D float noiseLerp(float t, float a, float b) { return a + t * (b - a); }
YG_D float newPerlinGrad(int hash, float x, float y, float z)
{	// noise_generator.h:50-56: the low four bits pick one of twelve edge directions
	const int h = hash & 15;
	const float u = h < 8 ? x : y, v = h < 4 ? y : ((h == 12 || h == 14) ? x : z);
	return ((h & 1) == 0 ? u : -u) + ((h & 2) == 0 ? v : -v);
}
YG_D float newPerlinFade(float t) { return t * t * t * (t * (t * 6 - 15) + 10); }
YG_D float noiseNewPerlin(const NoiseTables &nt, V3 pt)
{	// noise_generator.cc:182-205
	float x = pt.x, y = pt.y, z = pt.z;
	float u = floorf(x), v = floorf(y), w = floorf(z);
	const int cx = ((int)u) & 255, cy = ((int)v) & 255, cz = ((int)w) & 255;
	x -= u; y -= v; z -= w;
	u = newPerlinFade(x); v = newPerlinFade(y); w = newPerlinFade(z);
	const unsigned char *hs = nt.hash;
	const int a = hs[cx] + cy, aa = hs[a] + cz, ab = hs[a + 1] + cz, b = hs[cx + 1] + cy, ba = hs[b] + cz, bb = hs[b + 1] + cz;
	const float lo = noiseLerp(v, noiseLerp(u, newPerlinGrad(hs[aa], x, y, z), newPerlinGrad(hs[ba], x - 1, y, z)),
	                              noiseLerp(u, newPerlinGrad(hs[ab], x, y - 1, z), newPerlinGrad(hs[bb], x - 1, y - 1, z)));
	const float hi = noiseLerp(v, noiseLerp(u, newPerlinGrad(hs[aa + 1], x, y, z - 1), newPerlinGrad(hs[ba + 1], x - 1, y, z - 1)),
	                              noiseLerp(u, newPerlinGrad(hs[ab + 1], x, y - 1, z - 1), newPerlinGrad(hs[bb + 1], x - 1, y - 1, z - 1)));
	const float nv = noiseLerp(w, lo, hi);
	return (float)(0.5 + 0.5 * (double)nv);
}
YG_D void stdPerlinSetup(float c, int &b0, int &b1, float &r0, float &r1)
{	// SETUP, noise_generator.cc:312-317
	const float t = (float)((double)c + 10000.0);
	b0 = ((int)t) & 255; b1 = (b0 + 1) & 255;
	r0 = t - (float)(int)t; r1 = (float)((double)r0 - 1.0);
}
YG_D float stdPerlinSurve(float t) { return (float)((double)(t * t) * (3.0 - 2.0 * (double)t)); }
YG_D float stdPerlinAt(const float *q, float rx, float ry, float rz) { return rx * q[0] + ry * q[1] + rz * q[2]; }
YG_D float noiseStdPerlin(const NoiseTables &nt, V3 pt)
{	// noise_generator.cc:323-374
	int bx0, bx1, by0, by1, bz0, bz1;
	float rx0, rx1, ry0, ry1, rz0, rz1;
	stdPerlinSetup(pt.x, bx0, bx1, rx0, rx1); stdPerlinSetup(pt.y, by0, by1, ry0, ry1); stdPerlinSetup(pt.z, bz0, bz1, rz0, rz1);
	const unsigned char *pp = nt.stdp_p;
	const int i = pp[bx0], j = pp[bx1];
	const int b00 = pp[i + by0], b10 = pp[j + by0], b01 = pp[i + by1], b11 = pp[j + by1];
	const float sx = stdPerlinSurve(rx0), sy = stdPerlinSurve(ry0), sz = stdPerlinSurve(rz0);
	const float *g = nt.stdp_g;
	float u = stdPerlinAt(g + 3 * (b00 + bz0), rx0, ry0, rz0), v = stdPerlinAt(g + 3 * (b10 + bz0), rx1, ry0, rz0);
	float a = noiseLerp(sx, u, v);
	u = stdPerlinAt(g + 3 * (b01 + bz0), rx0, ry1, rz0); v = stdPerlinAt(g + 3 * (b11 + bz0), rx1, ry1, rz0);
	float b = noiseLerp(sx, u, v);
	const float c = noiseLerp(sy, a, b);
	u = stdPerlinAt(g + 3 * (b00 + bz1), rx0, ry0, rz1); v = stdPerlinAt(g + 3 * (b10 + bz1), rx1, ry0, rz1);
	a = noiseLerp(sx, u, v);
	u = stdPerlinAt(g + 3 * (b01 + bz1), rx0, ry1, rz1); v = stdPerlinAt(g + 3 * (b11 + bz1), rx1, ry1, rz1);
	b = noiseLerp(sx, u, v);
	const float d = noiseLerp(sy, a, b);
	return (float)(0.5 + 0.75 * (double)noiseLerp(sz, c, d));
}
YG_D float blenderFall(float c, float o, double sign) { return (float)((1.0 - 3.0 * (double)c) + sign * ((2.0 * (double)c) * (double)o)); }
YG_D float noiseBlender(const NoiseTables &nt, V3 pt)
{	// noise_generator.cc:411-481: eight lattice corners, cubic fall-off per axis, gradients from hashvectf
	const float x = pt.x, y = pt.y, z = pt.z;
	const int ix = (int)floorf(x), iy = (int)floorf(y), iz = (int)floorf(z);
	const float ox = x - (float)ix, oy = y - (float)iy, oz = z - (float)iz;
	const float jx = ox - 1, jy = oy - 1, jz = oz - 1;
	const float fx0 = blenderFall(ox * ox, ox, 1.0), fy0 = blenderFall(oy * oy, oy, 1.0), fz0 = blenderFall(oz * oz, oz, 1.0);
	const float fx1 = blenderFall(jx * jx, jx, -1.0), fy1 = blenderFall(jy * jy, jy, -1.0), fz1 = blenderFall(jz * jz, jz, -1.0);
	const unsigned char *hs = nt.hash;
	const int b00 = hs[hs[ix & 255] + (iy & 255)], b10 = hs[hs[(ix + 1) & 255] + (iy & 255)];
	const int b01 = hs[hs[ix & 255] + ((iy + 1) & 255)], b11 = hs[hs[(ix + 1) & 255] + ((iy + 1) & 255)];
	const int b20 = iz & 255, b21 = (iz + 1) & 255;
	float n = 0.5f;
	const float *h;
#define YG_BN(wt, bz, bxy, dx, dy, dz) h = nt.hashvectf + 3 * hs[(bz) + (bxy)]; n += (wt) * (h[0] * (dx) + h[1] * (dy) + h[2] * (dz));
	YG_BN(fx0 * fy0 * fz0, b20, b00, ox, oy, oz)
	YG_BN(fx0 * fy0 * fz1, b21, b00, ox, oy, jz)
	YG_BN(fx0 * fy1 * fz0, b20, b01, ox, jy, oz)
	YG_BN(fx0 * fy1 * fz1, b21, b01, ox, jy, jz)
	YG_BN(fx1 * fy0 * fz0, b20, b10, jx, oy, oz)
	YG_BN(fx1 * fy0 * fz1, b21, b10, jx, oy, jz)
	YG_BN(fx1 * fy1 * fz0, b20, b11, jx, jy, oz)
	YG_BN(fx1 * fy1 * fz1, b21, b11, jx, jy, jz)
#undef YG_BN
	if(n < 0.f) n = 0.f; else if(n > 1.f) n = 1.f;
	return n;
}
YG_D float noiseVoronoi(const NoiseTables &nt, V3 pt, int kind)
{	// VoronoiNoiseGenerator::getFeatures + operator() noise_generator.cc:531-600 with the factory's defaults (DistReal): the four nearest of the
	// 27 jittered feature points around the cell
	const float x = pt.x, y = pt.y, z = pt.z;
	const int xi = (int)floorf(x), yi = (int)floorf(y), zi = (int)floorf(z);
	float da0 = 1e10f, da1 = 1e10f, da2 = 1e10f, da3 = 1e10f;
	const unsigned char *hs = nt.hash;
	for(int xx = xi - 1; xx <= xi + 1; ++xx)
		for(int yy = yi - 1; yy <= yi + 1; ++yy)
			for(int zz = zi - 1; zz <= zi + 1; ++zz)
			{
				const float *p = nt.hashpntf + 3 * hs[(hs[(hs[zz & 255] + yy) & 255] + xx) & 255];
				const float xd = x - (p[0] + (float)xx), yd = y - (p[1] + (float)yy), zd = z - (p[2] + (float)zz);
				const float d = fSqrt(xd * xd + yd * yd + zd * zd);
				if(d < da0) { da3 = da2; da2 = da1; da1 = da0; da0 = d; }
				else if(d < da1) { da3 = da2; da2 = da1; da1 = d; }
				else if(d < da2) { da3 = da2; da2 = d; }
				else if(d < da3) da3 = d;
			}
	switch(kind)
	{
		case YAFGPU_NOISE_VORONOI_F2: return da1;
		case YAFGPU_NOISE_VORONOI_F3: return da2;
		case YAFGPU_NOISE_VORONOI_F4: return da3;
		case YAFGPU_NOISE_VORONOI_F2F1: return da1 - da0;
		case YAFGPU_NOISE_VORONOI_CRACKLE: { const float t = (float)(10.0 * (double)(da1 - da0)); return ((double)t > 1.0) ? 1.f : t; }
		default: return da0;
	}
}
YG_D float noiseCell(V3 pt)
{	// noise_generator.cc:606-614: an integer hash of the cell, scaled by 2^-32 in double
	const int xi = (int)floorf(pt.x), yi = (int)floorf(pt.y), zi = (int)floorf(pt.z);
	unsigned n = (unsigned)xi + (unsigned)yi * 1301u + (unsigned)zi * 314159u;
	n ^= (n << 13);
	return (float)((double)(float)(n * (n * n * 15731u + 789221u) + 1376312589u) / 4294967296.0);
}
YG_DN float noiseEval(const NoiseTables &nt, int type, V3 p)
{
	switch(type)
	{
		case YAFGPU_NOISE_STDPERLIN: return noiseStdPerlin(nt, p);
		case YAFGPU_NOISE_BLENDER: return noiseBlender(nt, p);
		case YAFGPU_NOISE_CELL: return noiseCell(p);
		case YAFGPU_NOISE_NEWPERLIN: return noiseNewPerlin(nt, p);
		default: return noiseVoronoi(nt, p, type);
	}
}
// turbulence__ noise_generator.cc:799-811: depth + 1 octaves, half amplitude and double frequency each; only the Blender basis offsets the point
YG_D float noiseTurbulence(const TexDev &t, V3 pt)
{
	float amp = 1.f, sum = 0.f;
	V3 tp = (t.noise_type == YAFGPU_NOISE_BLENDER ? pt + mk(1.f, 1.f, 1.f) : pt) * t.size;
	for(int i = 0; i <= t.depth; ++i, amp *= 0.5f, tp = tp * 2.f)
	{
		float val = noiseEval(*t.noise, t.noise_type, tp);
		if(t.hard) val = (float)fabs(2.0 * (double)val - 1.0);
		sum += amp * val;
	}
	return sum * ((float)(1 << t.depth) / (float)((1 << (t.depth + 1)) - 1));
}
// marble / wood wave shapes, texture_basic.cc:190-206,268-291
YG_D float noiseWave(int shape, float w)
{
	switch(shape)
	{
		case YAFGPU_WAVE_SAW: w *= (float)(0.5 * 0.318309886183790671538); w -= floorf(w); return w;
		case YAFGPU_WAVE_TRI: w *= (float)(0.5 * 0.318309886183790671538); return fabsf(2.f * (w - floorf(w)) - 1.f);
		default: return 0.5f + 0.5f * fSin(w);
	}
}
// VoronoiNoiseGenerator distance metrics, noise_generator.cc:377-409; setDistM (:487-521) maps DistManhattan to the SQUARED metric, as the reference does
YG_D float voronoiDistance(int metric, float x, float y, float z, float e)
{
	switch(metric)
	{
		case YAFGPU_DIST_SQUARED: case YAFGPU_DIST_MANHATTAN: return x * x + y * y + z * z;
		case YAFGPU_DIST_CHEBYCHEV: { x = fabsf(x); y = fabsf(y); z = fabsf(z); const float t = (x > y) ? x : y; return (z > t) ? z : t; }
		case YAFGPU_DIST_MINKOVSKY_HALF: { const float d = fSqrt(fabsf(x)) + fSqrt(fabsf(y)) + fSqrt(fabsf(z)); return d * d; }
		case YAFGPU_DIST_MINKOVSKY_FOUR: { x *= x; y *= y; z *= z; return fSqrt(fSqrt(x * x + y * y + z * z)); }
		case YAFGPU_DIST_MINKOVSKY: return fPow(fPow(fabsf(x), e) + fPow(fabsf(y), e) + fPow(fabsf(z), e), 1.f / e);
		default: return fSqrt(x * x + y * y + z * z);
	}
}
// VoronoiNoiseGenerator::getFeatures noise_generator.cc:531-577 with the feature points kept
YG_DN void voronoiFeatures(const NoiseTables &nt, int metric, float e, V3 pt, float *da, V3 *pa)
{
	const float x = pt.x, y = pt.y, z = pt.z;
	const int xi = (int)floorf(x), yi = (int)floorf(y), zi = (int)floorf(z);
	da[0] = da[1] = da[2] = da[3] = 1e10f;
	pa[0] = pa[1] = pa[2] = pa[3] = mk(0.f);
	const unsigned char *hs = nt.hash;
	for(int xx = xi - 1; xx <= xi + 1; ++xx)
		for(int yy = yi - 1; yy <= yi + 1; ++yy)
			for(int zz = zi - 1; zz <= zi + 1; ++zz)
			{
				const float *p = nt.hashpntf + 3 * hs[(hs[(hs[zz & 255] + yy) & 255] + xx) & 255];
				const V3 f = mk(p[0] + (float)xx, p[1] + (float)yy, p[2] + (float)zz);
				const float d = voronoiDistance(metric, x - f.x, y - f.y, z - f.z, e);
				if(d < da[0]) { da[3] = da[2]; da[2] = da[1]; da[1] = da[0]; da[0] = d; pa[3] = pa[2]; pa[2] = pa[1]; pa[1] = pa[0]; pa[0] = f; }
				else if(d < da[1]) { da[3] = da[2]; da[2] = da[1]; da[1] = d; pa[3] = pa[2]; pa[2] = pa[1]; pa[1] = f; }
				else if(d < da[2]) { da[3] = da[2]; da[2] = d; pa[3] = pa[2]; pa[2] = f; }
				else if(d < da[3]) { da[3] = d; pa[3] = f; }
			}
}
// cellNoiseColor__ noise_generator.cc:789-796
YG_D Rgba4 cellNoiseColor(const NoiseTables &nt, V3 pt)
{
	const int xi = (int)floorf(pt.x), yi = (int)floorf(pt.y), zi = (int)floorf(pt.z);
	const unsigned char *hs = nt.hash;
	const float *cl = nt.hashpntf + 3 * hs[(hs[(hs[zi & 255] + yi) & 255] + xi) & 255];
	return mk4(cl[0], cl[1], cl[2], 1.f);
}
// VoronoiTexture::getFloat / getColor texture_basic.cc:422-454 (no colour ramp)
YG_D float voronoiIntensity(const TexDev &t, const float *da)
{
	return t.iscale * fabsf(t.vw[0] * da[0] + t.vw[1] * da[1] + t.vw[2] * da[2] + t.vw[3] * da[3]);
}
YG_D Rgba4 voronoiTexColor(const TexDev &t, V3 p)
{
	float da[4]; V3 pa[4];
	voronoiFeatures(*t.noise, t.dist_metric, t.mk_exp, p * t.size, da, pa);
	const float inte = voronoiIntensity(t, da);
	if(t.n_ramp) return texAdjustColor(t, rampColor(t, inte));
	if(!t.color_type) return texAdjustColor(t, mk4(inte, inte, inte, inte));
	Rgba4 col = mk4(0.f, 0.f, 0.f, 0.f);   // Rgba(0.0): alpha 0 as well
	for(int k = 0; k < 4; ++k) col = col + t.vaw[k] * cellNoiseColor(*t.noise, pa[k]);
	if(t.color_type >= 2)
	{
		float t1 = (float)((double)(da[1] - da[0]) * 10.0);
		if(t1 > 1) t1 = 1;
		if(t.color_type == 3) t1 *= inte; else t1 *= t.iscale;
		col = t1 * col;
	}
	else col = t.iscale * col;
	return texAdjust(t, col);
}
// getSignedNoise__ noise_generator.h:263-266
YG_D float noiseSigned(const NoiseTables &nt, int type, V3 p) { return 2.f * noiseEval(nt, type, p) - 1.f; }
// the five Musgrave fractals, noise_generator.cc:631-784
YG_D float musgraveEval(const TexDev &t, V3 pt)
{
	const NoiseTables &nt = *t.noise;
	const int ng = t.noise_type, octs = (int)t.mus_oct;
	const float lac = t.mus_lac, offs = t.mus_offset, gain = t.mus_gain;
	const float pw_hl = fPow(lac, -t.mus_h);
	const float rmd = t.mus_oct - floorf(t.mus_oct);
	V3 tp = pt;
	switch(t.mus_type)
	{
		case YAFGPU_MUSGRAVE_MULTIFRACTAL:
		{
			float value = 1, pwr = 1;
			for(int i = 0; i < octs; ++i) { value *= (pwr * noiseSigned(nt, ng, tp) + 1.f); pwr *= pw_hl; tp = tp * lac; }
			if(rmd != 0.f) value *= (rmd * noiseSigned(nt, ng, tp) * pwr + 1.f);
			return value;
		}
		case YAFGPU_MUSGRAVE_HETEROTERRAIN:
		{
			float pwr = pw_hl;
			float value = offs + noiseSigned(nt, ng, tp);
			tp = tp * lac;
			for(int i = 1; i < octs; ++i)
			{
				const float increment = (noiseSigned(nt, ng, tp) + offs) * pwr * value;
				value += increment; pwr *= pw_hl; tp = tp * lac;
			}
			if(rmd != 0.f) { const float increment = (noiseSigned(nt, ng, tp) + offs) * pwr * value; value += rmd * increment; }
			return value;
		}
		case YAFGPU_MUSGRAVE_HYBRIDMF:
		{
			float pwr = pw_hl;
			float result = noiseSigned(nt, ng, tp) + offs;
			float weight = gain * result;
			tp = tp * lac;
			for(int i = 1; (weight > 0.001f) && (i < octs); ++i)
			{
				if(weight > 1.f) weight = 1.f;
				const float signal = (noiseSigned(nt, ng, tp) + offs) * pwr;
				pwr *= pw_hl;
				result += weight * signal;
				weight *= gain * signal;
				tp = tp * lac;
			}
			if(rmd != 0.f) result += rmd * ((noiseSigned(nt, ng, tp) + offs) * pwr);
			return result;
		}
		case YAFGPU_MUSGRAVE_RIDGEDMF:
		{
			float pwr = pw_hl;
			float signal = offs - fabsf(noiseSigned(nt, ng, tp));
			signal *= signal;
			float result = signal, weight = 1.f;
			for(int i = 1; i < octs; ++i)
			{
				tp = tp * lac;
				weight = signal * gain;
				if(weight > 1.f) weight = 1.f; else if(weight < 0.f) weight = 0.f;
				signal = offs - fabsf(noiseSigned(nt, ng, tp));
				signal *= signal;
				signal *= weight;
				result += signal * pwr;
				pwr *= pw_hl;
			}
			return result;
		}
		default:
		{	// fBm
			float value = 0, pwr = 1;
			for(int i = 0; i < octs; ++i) { value += noiseSigned(nt, ng, tp) * pwr; pwr *= pw_hl; tp = tp * lac; }
			if(rmd != 0.f) value += rmd * noiseSigned(nt, ng, tp) * pwr;
			return value;
		}
	}
}

// CloudsTexture / MarbleTexture / WoodTexture ::getFloat
YG_D float noiseTexFloat(const TexDev &t, V3 p)
{
	if(t.type == YAFGPU_TEX_VORONOI)
	{
		float da[4]; V3 pa[4];
		voronoiFeatures(*t.noise, t.dist_metric, t.mk_exp, p * t.size, da, pa);
		return texAdjustFloat(t, voronoiIntensity(t, da));
	}
	if(t.type == YAFGPU_TEX_MUSGRAVE) return texAdjustFloat(t, t.iscale * musgraveEval(t, p * t.size));
	if(t.type == YAFGPU_TEX_DISTORTED_NOISE)
	{	// texture_basic.cc:626-633: a noise vector from the first basis displaces the lookup into the second
		const V3 ofs = mk(13.5f, 13.5f, 13.5f), tp = p * t.size;
		const V3 rv = mk(noiseSigned(*t.noise, t.noise_type, tp + ofs), noiseSigned(*t.noise, t.noise_type, tp), noiseSigned(*t.noise, t.noise_type, tp - ofs));
		return texAdjustFloat(t, noiseSigned(*t.noise, t.noise_type2, tp + rv * t.distort));
	}
	if(t.type == YAFGPU_TEX_CLOUDS)
	{
		float v = noiseTurbulence(t, p);
		if(t.bias)
		{
			v *= v;
			if(t.bias == 1) return -v;   // returned without the adjustments, as the reference does (:118)
		}
		return texAdjustFloat(t, v);
	}
	if(t.type == YAFGPU_TEX_MARBLE)
	{	// `(x + y + z) * 5.0 + turb * turbulence` is a double sum narrowed into w
		const double tb = ((double)t.turb == 0.0) ? 0.0 : (double)(t.turb * noiseTurbulence(t, p));
		const float w = (float)((double)(p.x + p.y + p.z) * 5.0 + tb);
		return texAdjustFloat(t, fPow(noiseWave(t.shape, w), t.sharpness));
	}
	float w;
	if(t.rings) w = (float)((double)fSqrt(p.x * p.x + p.y * p.y + p.z * p.z) * 20.0);
	else w = (float)((double)(p.x + p.y + p.z) * 10.0);
	w = (float)((double)w + (((double)t.turb == 0.0) ? 0.0 : (double)(t.turb * noiseTurbulence(t, p))));
	return texAdjustFloat(t, noiseWave(t.shape, w));
}

// BlendTexture::getFloat texture_basic.cc:698-751 (its factory leaves the adjustments unset)
YG_D float blendTexFloat(const TexDev &t, V3 p)
{
	float blend = 0.f;
	const float coord_1 = t.flip ? p.y : p.x, coord_2 = t.flip ? p.x : p.y;
	switch(t.blend_type)
	{
		case YAFGPU_TEXBLEND_QUADRATIC:
			blend = 0.5f * (coord_1 + 1.f);
			if(blend < 0.f) blend = 0.f; else blend *= blend;
			break;
		case YAFGPU_TEXBLEND_EASING:
			blend = 0.5f * (coord_1 + 1.f);
			if(blend <= 0.f) blend = 0.f;
			else if(blend >= 1.f) blend = 1.f;
			else blend = (3.f * blend * blend - 2.f * blend * blend * blend);
			break;
		case YAFGPU_TEXBLEND_DIAGONAL: blend = 0.25f * (2.f + coord_1 + coord_2); break;
		case YAFGPU_TEXBLEND_SPHERICAL: case YAFGPU_TEXBLEND_QUADRATIC_SPHERE:
			blend = 1.f - fSqrt(coord_1 * coord_1 + coord_2 * coord_2 + p.z * p.z);
			if(blend < 0.f) blend = 0.f;
			if(t.blend_type == YAFGPU_TEXBLEND_QUADRATIC_SPHERE) blend *= blend;
			break;
		case YAFGPU_TEXBLEND_RADIAL: blend = (atan2f(coord_2, coord_1) / (float)(2.f * 3.14159265358979323846) + 0.5f); break;
		default: blend = 0.5f * (coord_1 + 1.f); break;
	}
	return fmaxf(0.f, fminf(blend, 1.f));
}

// ImageTexture::getColor(Point3) :75-87; BlendTexture::getColor texture_basic.cc:753-757 (no colour ramp); RgbCubeTexture::getColor :347-354
YG_D Rgba4 texGetColor(const TexDev &t, V3 p)
{
	if(t.type == YAFGPU_TEX_BLEND) { const float f = blendTexFloat(t, p); return t.n_ramp ? rampColor(t, f) : mk4(f, f, f, 1.f); }
	if(t.type >= YAFGPU_TEX_CLOUDS)
	{	// color_1_ + getFloat(p) * (color_2_ - color_1_), then applyColorAdjustments (texture_basic.cc:122-126,208-212,293-297)
		if(t.type == YAFGPU_TEX_VORONOI) return voronoiTexColor(t, p);
		const float f = noiseTexFloat(t, p);
		if(t.n_ramp) return texAdjustColor(t, rampColor(t, f));
		const Rgba4 c = mk4(t.color1[0] + f * (t.color2[0] - t.color1[0]), t.color1[1] + f * (t.color2[1] - t.color1[1]), t.color1[2] + f * (t.color2[2] - t.color1[2]), 1.f);
		return texAdjustColor(t, c);
	}
	if(t.type == YAFGPU_TEX_RGB_CUBE)
	{
		Rgba4 c = mk4(fmaxf(0.f, fminf(1.f, p.x)), fmaxf(0.f, fminf(1.f, p.y)), fmaxf(0.f, fminf(1.f, p.z)), 1.f);
		return t.adj_set ? texAdjust(t, c) : c;
	}
	float x = p.x, y = -p.y, z = p.z;
	if(imgDoMapping(t, x, y, z)) return mk4(0.f, 0.f, 0.f, 0.f);
	return texAdjust(t, imgInterpolate(t, x, y));
}
// ImageTexture::getRawColor(Point3) :89-103
YG_D Rgba4 texGetRawColor(const TexDev &t, V3 p)
{
	if(t.type != YAFGPU_TEX_IMAGE) return texGetColor(t, p); // Texture::getRawColor texture.h:48
	return colorSpaceFromLinear(texGetColor(t, p), t.color_space, t.gamma);
}
// Texture::getFloat texture.h:51; RgbCubeTexture::getFloat texture_basic.cc:356-361 (Rgb::energy = (r + g + b) * 0.333333f)
YG_D float texGetFloat(const TexDev &t, V3 p)
{
	if(t.type == YAFGPU_TEX_BLEND) return blendTexFloat(t, p);
	if(t.type >= YAFGPU_TEX_CLOUDS) return noiseTexFloat(t, p);
	if(t.type == YAFGPU_TEX_RGB_CUBE)
		return texAdjustFloat(t, (fmaxf(0.f, fminf(1.f, p.x)) + fmaxf(0.f, fminf(1.f, p.y)) + fmaxf(0.f, fminf(1.f, p.z))) * 0.333333f);
	return texAdjustFloat(t, col2Bri(texGetRawColor(t, p)));
}

// ---- TextureMapperNode ------------------------------------------------------------------------------------------------------
// doMapping shader_node_basic.cc:122-152 with tubemap__/spheremap__/cubemap__ :57-113
YG_D V3 mapperDoMapping(const NodeDev &nd, V3 p, V3 n)
{
	V3 t = p;
	if(nd.coords == YAFGPU_TEXCO_UV) t = mk(2.0f * t.x - 1.0f, 2.0f * t.y - 1.0f, t.z);
	const float texmap[4] = {0.f, t.x, t.y, t.z};
	t = mk(texmap[nd.map_x], texmap[nd.map_y], texmap[nd.map_z]);
	switch(nd.proj)
	{
		case YAFGPU_TEXPROJ_TUBE:
		{
			V3 res;
			res.y = t.z;
			const float d = t.x * t.x + t.y * t.y;
			if(d > 0)
			{
				res.z = (float)(1.0 / (double)fSqrt(d));
				res.x = (float)(-atan2((double)t.x, (double)t.y) * YG_M_1_PI); // unqualified atan2/floor on floats resolve to the C double functions
			}
			else res.x = res.z = 0;
			t = res;
			break;
		}
		case YAFGPU_TEXPROJ_SPHERE:
		{
			V3 res = mk(0.f);
			const float d = (t.x * t.x + t.y * t.y) + t.z * t.z;
			if(d > 0)
			{
				res.z = fSqrt(d);
				if((t.x != 0) && (t.y != 0)) res.x = (float)(-atan2((double)t.x, (double)t.y) * YG_M_1_PI);
				res.y = (float)(1.0 - 2.0 * ((double)fAcos(t.z / res.z) * YG_M_1_PI));
			}
			t = res;
			break;
		}
		case YAFGPU_TEXPROJ_CUBE:
		{
			int axis;
			if(fabsf(n.z) >= fabsf(n.x) && fabsf(n.z) >= fabsf(n.y)) axis = 2;
			else if(fabsf(n.y) >= fabsf(n.x) && fabsf(n.y) >= fabsf(n.z)) axis = 1;
			else axis = 0;
			if(axis == 0) t = mk(t.y, t.z, t.x);
			else if(axis == 1) t = mk(t.x, t.z, t.y);
			break;
		}
		default: break;
	}
	return mk(t.x * nd.scale[0], t.y * nd.scale[1], t.z * nd.scale[2]) + mk(nd.offset[0], nd.offset[1], nd.offset[2]);
}
// getCoords :159-186 (the "not implemented yet" modes map to global as there)
YG_D void mapperGetCoords(const CameraDev &cam, const NodeDev &nd, const SurfPt &sp, const SurfTex &sx, V3 &texpt, V3 &ng)
{
	switch(nd.coords)
	{
		case YAFGPU_TEXCO_UV: texpt = mk(sx.u, sx.v, 0.f); ng = sp.ng; break;
		case YAFGPU_TEXCO_ORCO: texpt = sx.orco_p; ng = sx.orco_ng; break;
		case YAFGPU_TEXCO_TRANSFORMED:
		{
			const float *m = nd.mtx;
			texpt = mk(((m[0] * sp.p.x + m[1] * sp.p.y) + m[2] * sp.p.z) + m[3], ((m[4] * sp.p.x + m[5] * sp.p.y) + m[6] * sp.p.z) + m[7], ((m[8] * sp.p.x + m[9] * sp.p.y) + m[10] * sp.p.z) + m[11]);
			ng = mk((m[0] * sp.ng.x + m[1] * sp.ng.y) + m[2] * sp.ng.z, (m[4] * sp.ng.x + m[5] * sp.ng.y) + m[6] * sp.ng.z, (m[8] * sp.ng.x + m[9] * sp.ng.y) + m[10] * sp.ng.z);
			break;
		}
		case YAFGPU_TEXCO_WINDOW:
		{	// Camera::screenproject of the five camera types
			const V3 cx = ldv(cam.axis_x), cy = ldv(cam.axis_y), cz = ldv(cam.axis_z);
			V3 dir = sp.p - ldv(cam.pos);   // orthographic: cam.pos holds pos_ (the corner of the image plane), as screenproject uses it
			ng = sp.ng;
			if(cam.type == YAFGPU_CAM_ORTHOGRAPHIC)
			{	// camera_orthographic.cc:65-80
				const float dz = dot(cz, dir);
				const V3 proj = dir - dz * cz;
				texpt = mk(2 * (dot(proj, cx) / cam.ortho_scale) - 1.0f, dot(-2.f * proj, cy) / (cam.aspect_ratio * cam.ortho_scale) + 1.0f, 0.f);
			}
			else if(cam.type == YAFGPU_CAM_ANGULAR || cam.type == YAFGPU_CAM_EQUIRECTANGULAR)
			{	// camera_angular.cc:123-140, camera_equirectangular.cc:92-109: the same (FIXME-marked) formula, the quotient formed in double
				dir = normalize(dir);
				const float dx = dot(cx, dir), dy = dot(cy, dir), dz = dot(cz, dir);
				texpt = mk((float)((double)-dx / (4.0 * YG_M_PI * (double)dz)), (float)((double)-dy / (4.0 * YG_M_PI * (double)dz)), 0.f);
			}
			else if(cam.type == YAFGPU_CAM_ARCHITECT)
			{	// camera_architect.cc:68-91: vertical axis fixed to world z, focal distance corrected by the tilt
				const V3 camy = mk(0.f, 0.f, 1.f), camz = cross(camy, cx), camx = cross(camz, camy);
				const float dx = dot(dir, camx), dy = dot(dir, cy), dz = dot(dir, camz);
				const float fod = dot(cam.focal * camy, cy) / dot(camx, cx);
				texpt = mk(2 * dx * fod / dz, 2 * dy * cam.focal / (dz * cam.aspect_ratio), 0.f);
			}
			else
			{	// camera_perspective.cc:158-173
				const float dx = dot(dir, cx), dy = dot(dir, cy), dz = dot(dir, cz);
				texpt = mk(2.0f * dx * cam.focal / dz, -2.0f * dy * cam.focal / (dz * cam.aspect_ratio), 0.f);
			}
			break;
		}
		case YAFGPU_TEXCO_NORMAL: // Camera::getAxis; sp.n is the (possibly bumped) shading normal at this point of initBsdf
			texpt = mk(dot(sp.n, ldv(cam.axis_x)), -dot(sp.n, ldv(cam.axis_y)), 0.f); ng = sp.ng;
			break;
		default: texpt = sp.p; ng = sp.ng; break;
	}
}
// eval :189-231 (no mipmap interpolation types on the device)
YG_D NodeRes mapperEval(const SceneDev &sc, const NodeDev &nd, const SurfPt &sp, const SurfTex &sx)
{
	V3 texpt, ng;
	mapperGetCoords(sc.cam, nd, sp, sx, texpt, ng);
	texpt = mapperDoMapping(nd, texpt, ng);
	const TexDev &t = sc.textures[nd.tex];
	NodeRes r;
	r.col = texGetColor(t, texpt);
	r.f = nd.do_scalar ? texGetFloat(t, texpt) : 0.f;
	return r;
}
// evalDerivative :241-342
YG_D NodeRes mapperEvalDerivative(const SceneDev &sc, const NodeDev &nd, const SurfPt &sp, const SurfTex &sx)
{
	V3 texpt, ng;
	float du = 0.0f, dv = 0.0f;
	mapperGetCoords(sc.cam, nd, sp, sx, texpt, ng);
	const TexDev &t = sc.textures[nd.tex];
	if(nd.discrete && sx.has_uv && nd.coords == YAFGPU_TEXCO_UV)
	{
		texpt = mapperDoMapping(nd, texpt, ng);
		V3 norm;
		if(t.normalmap)
		{
			const Rgba4 color = texGetRawColor(t, texpt);
			norm = mk(2.f * color.r - 1.f, 2.f * color.g - 1.f, 2.f * color.b - 1.f);
			du = dot(norm, sx.ds_du); dv = dot(norm, sx.ds_dv);
		}
		else
		{
			const V3 p_du = mk(nd.d_u, 0.f, 0.f), p_dv = mk(0.f, nd.d_v, 0.f);
			const float dfdu = (texGetFloat(t, texpt - p_du) - texGetFloat(t, texpt + p_du)) / nd.d_u;
			const float dfdv = (texGetFloat(t, texpt - p_dv) - texGetFloat(t, texpt + p_dv)) / nd.d_v;
			V3 vec_u = sx.ds_du, vec_v = sx.ds_dv;
			vec_u.z = dfdu; vec_v.z = dfdv;
			norm = cross(vec_u, vec_v);
		}
		norm = normalize(norm);
		if(fabsf(norm.z) > 1e-30f)
		{
			const float nf = (float)(1.0 / (double)norm.z * (double)nd.bump_str);
			du = norm.x * nf; dv = norm.y * nf;
		}
		else du = dv = 0.f;
	}
	else if(t.normalmap)
	{
		texpt = mapperDoMapping(nd, texpt, ng);
		const Rgba4 color = texGetRawColor(t, texpt);
		V3 norm = mk(2.f * color.r - 1.f, 2.f * color.g - 1.f, 2.f * color.b - 1.f);
		du = dot(norm, sx.ds_du); dv = dot(norm, sx.ds_dv);
		norm = normalize(norm);
		if(fabsf(norm.z) > 1e-30f)
		{
			const float nf = (float)(1.0 / (double)norm.z * (double)nd.bump_str);
			du = norm.x * nf; dv = norm.y * nf;
		}
		else du = dv = 0.f;
	}
	else
	{
		const V3 i0 = mapperDoMapping(nd, texpt - nd.d_u * sp.nu, ng), i1 = mapperDoMapping(nd, texpt + nd.d_u * sp.nu, ng);
		const V3 j0 = mapperDoMapping(nd, texpt - nd.d_v * sp.nv, ng), j1 = mapperDoMapping(nd, texpt + nd.d_v * sp.nv, ng);
		du = (texGetFloat(t, i0) - texGetFloat(t, i1)) / nd.d_u;
		dv = (texGetFloat(t, j0) - texGetFloat(t, j1)) / nd.d_v;
		du *= nd.bump_str; dv *= nd.bump_str;
		if(nd.coords != YAFGPU_TEXCO_UV) { du = -du; dv = -dv; }
	}
	NodeRes r; r.col = mk4(du, dv, 0.f, 0.f); r.f = 0.f;
	return r;
}

// ---- MixNode family :452-690 ------------------------------------------------------------------------------------------------
YG_D float ovl(float c1, float c2, float f1, float f2)
{	// OverlayNode: float * double-literal expressions run in double
	return (c1 < 0.5f) ? c1 * (f1 + 2.0f * f2 * c2) : (float)(1.0 - ((double)f1 + (double)(2.0f * f2) * (1.0 - (double)c2)) * (1.0 - (double)c1));
}
YG_D NodeRes mixEval(const NodeDev &nd, const NodeRes *stack)
{
	const float f2 = nd.fac >= 0 ? stack[nd.fac].f : nd.f1; // factor_ ? getScalar : cfactor_
	Rgba4 c1, c2; float s1, s2;
	if(nd.in1 >= 0) { c1 = stack[nd.in1].col; s1 = stack[nd.in1].f; } else { c1 = mk4(nd.col[0], nd.col[1], nd.col[2], nd.col[3]); s1 = 0.f; }
	if(nd.in2 >= 0) { c2 = stack[nd.in2].col; s2 = stack[nd.in2].f; } else { c2 = mk4(nd.col2[0], nd.col2[1], nd.col2[2], nd.col2[3]); s2 = 0.f; }
	const float f1 = 1.f - f2;
	NodeRes r;
	switch(nd.mode)
	{
		case YAFGPU_MIX_ADD: r.col = c1 + f2 * c2; r.f = s1 + f2 * s2; break;
		case YAFGPU_MIX_MULT: r.col = c1 * (mk4(f1, f1, f1, f1) + f2 * c2); r.f = s1; break; // the scalar product is written to fin_2 and dropped (:541-543)
		case YAFGPU_MIX_SUB: r.col = c1 - f2 * c2; r.f = s1 - f2 * s2; break;
		case YAFGPU_MIX_SCREEN:
		{
			const Rgba4 one = mk4(1.f, 1.f, 1.f, 1.f);
			r.col = one - (mk4(f1, f1, f1, f1) + f2 * (one - c2)) * (one - c1); // Rgba(1.f) - ... ; (1.f - cin_2) is Rgba(1.f) - cin_2
			r.f = (float)(1.0 - (double)((f1 + f2 * (1.f - s2)) * (1.f - s1)));
			break;
		}
		case YAFGPU_MIX_DIFF:
			r.col = mk4(f1 * c1.r + f2 * fabsf(c1.r - c2.r), f1 * c1.g + f2 * fabsf(c1.g - c2.g), f1 * c1.b + f2 * fabsf(c1.b - c2.b), f1 * c1.a + f2 * fabsf(c1.a - c2.a));
			r.f = f1 * s1 + f2 * fabsf(s1 - s2);
			break;
		case YAFGPU_MIX_DARK:
		{
			c2 = c2 * f2;
			if(c2.r < c1.r) c1.r = c2.r; if(c2.g < c1.g) c1.g = c2.g; if(c2.b < c1.b) c1.b = c2.b; if(c2.a < c1.a) c1.a = c2.a;
			s2 *= f2; if(s2 < s1) s1 = s2;
			r.col = c1; r.f = s1;
			break;
		}
		case YAFGPU_MIX_LIGHT:
		{
			c2 = c2 * f2;
			if(c2.r > c1.r) c1.r = c2.r; if(c2.g > c1.g) c1.g = c2.g; if(c2.b > c1.b) c1.b = c2.b; if(c2.a > c1.a) c1.a = c2.a;
			s2 *= f2; if(s2 > s1) s1 = s2;
			r.col = c1; r.f = s1;
			break;
		}
		case YAFGPU_MIX_OVERLAY:
			r.col = mk4(ovl(c1.r, c2.r, f1, f2), ovl(c1.g, c2.g, f1, f2), ovl(c1.b, c2.b, f1, f2), ovl(c1.a, c2.a, f1, f2));
			r.f = ovl(s1, s2, f1, f2);
			break;
		default: // MnMix, and MnDiv which the factory maps to the plain mix node (:692-705)
			r.col = f1 * c1 + f2 * c2; r.f = f1 * s1 + f2 * s2;
			break;
	}
	return r;
}

// ---- LayerNode --------------------------------------------------------------------------------------------------------------
// textureRgbBlend__ shader_node.h:115-175 (Rgb, the alpha of the inputs plays no part)
YG_D Col texRgbBlend(Col tex, Col out, float fact, float facg, int mode)
{
	switch(mode)
	{
		case YAFGPU_MIX_MULT: fact *= facg; return (mkc(1.f - facg) + fact * tex) * out;
		case YAFGPU_MIX_SCREEN:
		{
			const Col white = mkc(1.0f);
			fact *= facg;
			const Col wt = mkc(white.r - tex.r, white.g - tex.g, white.b - tex.b), wo = mkc(white.r - out.r, white.g - out.g, white.b - out.b);
			const Col q = (mkc(1.f - facg) + fact * wt) * wo;
			return mkc(white.r - q.r, white.g - q.g, white.b - q.b);
		}
		case YAFGPU_MIX_SUB: fact = -fact;
		case YAFGPU_MIX_ADD: fact *= facg; return fact * tex + out;
		case YAFGPU_MIX_DIV:
		{
			fact *= facg;
			Col itex = tex;
			if(itex.r != 0.f) itex.r = 1.f / itex.r; if(itex.g != 0.f) itex.g = 1.f / itex.g; if(itex.b != 0.f) itex.b = 1.f / itex.b;
			return (1.f - fact) * out + (fact * out) * itex;
		}
		case YAFGPU_MIX_DIFF:
		{
			fact *= facg;
			const Col tmo = mkc(fabsf(tex.r - out.r), fabsf(tex.g - out.g), fabsf(tex.b - out.b));
			return (1.f - fact) * out + fact * tmo;
		}
		case YAFGPU_MIX_DARK:
		{
			fact *= facg;
			Col col = fact * tex;
			if(col.r > out.r) col.r = out.r; if(col.g > out.g) col.g = out.g; if(col.b > out.b) col.b = out.b;
			return col;
		}
		case YAFGPU_MIX_LIGHT:
		{
			fact *= facg;
			Col col = fact * tex;
			if(col.r < out.r) col.r = out.r; if(col.g < out.g) col.g = out.g; if(col.b < out.b) col.b = out.b;
			return col;
		}
		default: fact *= facg; return fact * tex + (1.f - fact) * out;
	}
}
// textureValueBlend__ shader_node.h:177-223 (flip = false)
YG_D float texValueBlend(float tex, float out, float fact, float facg, int mode)
{
	fact *= facg;
	float facm = 1.f - fact;
	switch(mode)
	{
		case YAFGPU_MIX_MULT: facm = 1.f - facg; return (facm + fact * tex) * out;
		case YAFGPU_MIX_SCREEN: facm = 1.f - facg; return 1.f - (facm + fact * (1.f - tex)) * (1.f - out);
		case YAFGPU_MIX_SUB: fact = -fact;
		case YAFGPU_MIX_ADD: return fact * tex + out;
		case YAFGPU_MIX_DIV: if(tex == 0.f) return 0.f; return facm * out + fact * out / tex;
		case YAFGPU_MIX_DIFF: return facm * out + fact * fabsf(tex - out);
		case YAFGPU_MIX_DARK: { const float col = fact * tex; return col < out ? col : out; }
		case YAFGPU_MIX_LIGHT: { const float col = fact * tex; return col > out ? col : out; }
		default: return fact * tex + facm * out;
	}
}
enum { TXF_RGBTOINT = 1, TXF_STENCIL = 2, TXF_NEGATIVE = 4, TXF_ALPHAMIX = 8 };
// LayerNode::eval shader_node_layer.cc:31-117
YG_D NodeRes layerEval(const NodeDev &nd, const NodeRes *stack)
{
	Rgba4 rcol, texcolor = mk4(0.f, 0.f, 0.f, 1.f); // Rgba texcolor; default-constructed: rgb 0, alpha 1
	float rval, tin = 0.f, ta = 1.f, stencil_tin;
	if(nd.in2 >= 0) { rcol = stack[nd.in2].col; rval = stack[nd.in2].f; }
	else { rcol = mk4(nd.col2[0], nd.col2[1], nd.col2[2], nd.col2[3]); rval = nd.f2; }
	stencil_tin = rcol.a;
	bool tex_rgb = nd.color_input != 0;
	if(nd.color_input) { texcolor = stack[nd.in1].col; ta = texcolor.a; }
	else tin = stack[nd.in1].f;
	if(nd.texflag & TXF_RGBTOINT) { tin = col2Bri(texcolor); tex_rgb = false; }
	if(nd.texflag & TXF_NEGATIVE)
	{
		if(tex_rgb) texcolor = mk4(1.f - texcolor.r, 1.f - texcolor.g, 1.f - texcolor.b, 1.f - texcolor.a);
		tin = 1.f - tin;
	}
	if(nd.texflag & TXF_STENCIL)
	{
		float fact;
		if(tex_rgb) { fact = ta; ta *= stencil_tin; stencil_tin *= fact; }
		else { fact = tin; tin *= stencil_tin; stencil_tin *= fact; }
	}
	if(nd.do_color)
	{
		if(!tex_rgb) texcolor = mk4(nd.col[0], nd.col[1], nd.col[2], nd.col[3]);
		else tin = ta;
		const float tin_tr = tin > 1.f ? 1.f : (tin < 0.f ? 0.f : tin);
		const Col c = texRgbBlend(rgbOf(texcolor), rgbOf(rcol), tin_tr, stencil_tin * nd.colfac, nd.mode);
		rcol = mk4(c.r, c.g, c.b, 1.f);
		clampRgb0(rcol);
	}
	if(nd.do_scalar_l)
	{
		if(tex_rgb)
		{
			if(nd.use_alpha) { tin = ta; if(nd.texflag & TXF_NEGATIVE) tin = 1.f - tin; }
			else tin = col2Bri(texcolor);
		}
		rval = texValueBlend(nd.f1, rval, tin, stencil_tin * nd.valfac, nd.mode);
		if(rval < 0.f) rval = 0.f;
	}
	rcol.a = stencil_tin;
	NodeRes r; r.col = rcol; r.f = rval;
	return r;
}
// LayerNode::evalDerivative :124-151
YG_D NodeRes layerEvalDerivative(const NodeDev &nd, const NodeRes *stack)
{
	float rdu = 0.f, rdv = 0.f, stencil_tin = 1.f;
	if(nd.in2 >= 0) { const Rgba4 u = stack[nd.in2].col; rdu = u.r; rdv = u.g; stencil_tin = u.a; }
	const Rgba4 tc = stack[nd.in1].col;
	float tdu = tc.r, tdv = tc.g;
	if(nd.texflag & TXF_NEGATIVE) { tdu = -tdu; tdv = -tdv; }
	rdu += tdu; rdv += tdv;
	NodeRes r; r.col = mk4(rdu, rdv, 0.f, stencil_tin); r.f = 0.f;
	return r;
}

// ---- the per-hit node pass ----------------------------------------------------------------------------------------------------
YG_D NodeRes nodeEval(const SceneDev &sc, const NodeDev &nd, const SurfPt &sp, const SurfTex &sx, const NodeRes *stack)
{
	switch(nd.type)
	{
		case YAFGPU_NODE_TEXTURE_MAPPER: return mapperEval(sc, nd, sp, sx);
		case YAFGPU_NODE_MIX: return mixEval(nd, stack);
		case YAFGPU_NODE_LAYER: return layerEval(nd, stack);
		default: { NodeRes r; r.col = mk4(nd.col[0], nd.col[1], nd.col[2], nd.col[3]); r.f = nd.f1; return r; } // ValueNode :426-434
	}
}
// ShaderNode::evalDerivative: only the texture mapper and the layer node override it; the base class leaves the stack slot untouched
// (shader_node.h) — a mix/value node inside a bump tree therefore yields whatever the slot held; it is written as zero here.
YG_D NodeRes nodeEvalDerivative(const SceneDev &sc, const NodeDev &nd, const SurfPt &sp, const SurfTex &sx, const NodeRes *stack)
{
	switch(nd.type)
	{
		case YAFGPU_NODE_TEXTURE_MAPPER: return mapperEvalDerivative(sc, nd, sp, sx);
		case YAFGPU_NODE_LAYER: return layerEvalDerivative(nd, stack);
		default: { NodeRes r; r.col = mk4(0.f, 0.f, 0.f, 0.f); r.f = 0.f; return r; }
	}
}

// What every node material's initBsdf() does first (e.g. material_shiny_diffuse.cc:164-183): evalBump with the bump node list, then the
// colour node list; afterwards every bound slot is folded into `ml`, the thread-local copy of the material record (see the header comment).
YG_DN void evalMaterialNodes(const SceneDev &sc, MatDev &ml, SurfPt &sp, int tri, float b1, float b2)
{
	SurfTex sx;
	getSurfaceTex(sc, tri, b1, b2, sp, sx);
	NodeRes stack[YG_MAX_NODES];
	const NodeDev *nodes = sc.nodes + ml.node_first;
	const int bump = ml.slot[YAFGPU_SLOT_BUMP];
	if(bump >= 0)
	{	// NodeMaterial::evalBump material_node.cc:132-139 + Material::applyBump material.cc:77-84
		for(int i = 0; i < ml.node_count; ++i) if(ml.bump_mask >> i & 1u) stack[i] = nodeEvalDerivative(sc, nodes[i], sp, sx, stack);
		const float du = stack[bump].col.r, dv = stack[bump].col.g;
		sp.nu = sp.nu + du * sp.n;
		sp.nv = sp.nv + dv * sp.n;
		sp.n = normalize(cross(sp.nu, sp.nv));
		sp.nu = normalize(sp.nu);
		sp.nv = normalize(cross(sp.n, sp.nu));
	}
	for(int i = 0; i < ml.node_count; ++i) if(ml.color_mask >> i & 1u) stack[i] = nodeEval(sc, nodes[i], sp, sx, stack);
#define YG_SLOT(S) ml.slot[YAFGPU_SLOT_##S]
	switch(ml.type)
	{
		case YAFGPU_MAT_SHINYDIFFUSE:
		{	// material_shiny_diffuse.cc: getComponents :98-118, eval :244-293, emit :295-307, sample :309-411, getSpecular :470-530
			if(YG_SLOT(DIFFUSE) >= 0)
			{
				const Rgba4 c = stack[YG_SLOT(DIFFUSE)].col;
				ml.diffuse_col[0] = c.r; ml.diffuse_col[1] = c.g; ml.diffuse_col[2] = c.b;
				ml.emit_col[0] = c.r * ml.emit_strength; ml.emit_col[1] = c.g * ml.emit_strength; ml.emit_col[2] = c.b * ml.emit_strength;
			}
			if(YG_SLOT(MIRROR_COLOR) >= 0) { const Rgba4 c = stack[YG_SLOT(MIRROR_COLOR)].col; ml.mirror_col[0] = c.r; ml.mirror_col[1] = c.g; ml.mirror_col[2] = c.b; }
			if(YG_SLOT(MIRROR) >= 0) ml.mirror_strength = stack[YG_SLOT(MIRROR)].f;
			if(YG_SLOT(TRANSPARENCY) >= 0) ml.transparency_strength = stack[YG_SLOT(TRANSPARENCY)].f;
			if(YG_SLOT(TRANSLUCENCY) >= 0) ml.translucency_strength = stack[YG_SLOT(TRANSLUCENCY)].f;
			if(YG_SLOT(DIFFUSE_REFL) >= 0) ml.diffuse_refl_mul = stack[YG_SLOT(DIFFUSE_REFL)].f;
			if(YG_SLOT(IOR) >= 0) { const float i = ml.ior + stack[YG_SLOT(IOR)].f; ml.ior_squared = i * i; }
			if(YG_SLOT(SIGMA_OREN) >= 0) { ml.oren_tex = 1; ml.oren_sigma = stack[YG_SLOT(SIGMA_OREN)].f; }
			break;
		}
		case YAFGPU_MAT_GLOSSY:
		case YAFGPU_MAT_COATED_GLOSSY:
		{	// material_glossy.cc:51-64,113-173,175-357; material_coated_glossy.cc:67-80,130-185,187-369,430-463
			if(YG_SLOT(DIFFUSE) >= 0) { const Rgba4 c = stack[YG_SLOT(DIFFUSE)].col; ml.diff_col[0] = c.r; ml.diff_col[1] = c.g; ml.diff_col[2] = c.b; }
			if(YG_SLOT(GLOSSY) >= 0) { const Rgba4 c = stack[YG_SLOT(GLOSSY)].col; ml.gloss_col[0] = c.r; ml.gloss_col[1] = c.g; ml.gloss_col[2] = c.b; }
			if(YG_SLOT(GLOSSY_REFLECT) >= 0)
			{	// initBsdf: m_glossy_ from the node, p_diffuse_ recomputed from it
				ml.gl_glossy = stack[YG_SLOT(GLOSSY_REFLECT)].f;
				ml.p_diffuse = fminf(0.6f, 1.f - (ml.gl_glossy / (ml.gl_glossy + (1.f - ml.gl_glossy) * ml.gl_diffuse)));
			}
			if(YG_SLOT(EXPONENT) >= 0) ml.gl_exponent = stack[YG_SLOT(EXPONENT)].f;
			if(YG_SLOT(DIFFUSE_REFL) >= 0) ml.diffuse_refl_mul = stack[YG_SLOT(DIFFUSE_REFL)].f;
			if(YG_SLOT(SIGMA_OREN) >= 0) { ml.oren_tex = 1; ml.oren_sigma = stack[YG_SLOT(SIGMA_OREN)].f; }
			if(ml.type == YAFGPU_MAT_COATED_GLOSSY)
			{
				if(YG_SLOT(MIRROR_COLOR) >= 0) { const Rgba4 c = stack[YG_SLOT(MIRROR_COLOR)].col; ml.gl_mirror_col[0] = c.r; ml.gl_mirror_col[1] = c.g; ml.gl_mirror_col[2] = c.b; }
				if(YG_SLOT(MIRROR) >= 0) ml.gl_mirror_strength = stack[YG_SLOT(MIRROR)].f;
				if(YG_SLOT(IOR) >= 0) ml.ior = ml.ior + stack[YG_SLOT(IOR)].f;
			}
			break;
		}
		case YAFGPU_MAT_GLASS:
		case YAFGPU_MAT_ROUGH_GLASS:
		{	// material_glass.cc:63-215,250-309; material_rough_glass.cc:60-170,172-290 (getTransparency reads the IOR node differently,
			// :223 / :298, but is only reached through transparent shadows, which are refused for node materials)
			if(YG_SLOT(FILTER_COLOR) >= 0) { const Rgba4 c = stack[YG_SLOT(FILTER_COLOR)].col; ml.filter_col[0] = c.r; ml.filter_col[1] = c.g; ml.filter_col[2] = c.b; }
			if(YG_SLOT(MIRROR_COLOR) >= 0) { const Rgba4 c = stack[YG_SLOT(MIRROR_COLOR)].col; ml.spec_refl_col[0] = c.r; ml.spec_refl_col[1] = c.g; ml.spec_refl_col[2] = c.b; }
			if(YG_SLOT(IOR) >= 0) ml.ior = ml.ior + stack[YG_SLOT(IOR)].f;
			if(ml.type == YAFGPU_MAT_ROUGH_GLASS && YG_SLOT(ROUGHNESS) >= 0) ml.alpha_rough = stack[YG_SLOT(ROUGHNESS)].f + 0.001f; // alpha_2 = alpha_texture^2 (:69-70)
			break;
		}
		case YAFGPU_MAT_MASK:
		{	// MaskMaterial::initBsdf material_mask.cc:40-51: every later call goes to the chosen child, so its record takes the mask's place;
			// what the integrator asks the *mask* object itself stays (visibility, shadow reception, additional depth, bias, sampling factor)
			const MatDev &ch = sc.materials[stack[YG_SLOT(BLEND)].f > ml.blend_val ? ml.mat2 : ml.mat1];
			const int vis = ml.visibility, rs = ml.receive_shadows, ad = ml.additional_depth, tbm = ml.transp_bias_mult;
			const float tbf = ml.transp_bias_factor, sf = ml.sampling_factor;
			const unsigned uni = ml.bsdf_flags; // MaskMaterial ctor: union of the children's flags
			ml = ch;
			ml.obj_flags = uni; ml.is_mask = 1;
			ml.absorb = 0; // MaskMaterial does not forward getVolumeHandler (the base class returns its own, unset, handlers)
			ml.visibility = vis; ml.receive_shadows = rs; ml.additional_depth = ad; ml.transp_bias_mult = tbm; ml.transp_bias_factor = tbf; ml.sampling_factor = sf;
			break;
		}
		case YAFGPU_MAT_BLEND: // BlendMaterial::getBlendVal material_blend.cc:57-75
			if(YG_SLOT(BLEND) >= 0) ml.blend_val = stack[YG_SLOT(BLEND)].f;
			break;
		default: break;
	}
#undef YG_SLOT
}

} // namespace yg
