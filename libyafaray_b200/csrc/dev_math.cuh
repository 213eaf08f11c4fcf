// Device (and host-mirror) math for the yafgpu hot path.
//
// Everything here restates, operation for operation, the scalar float code of the reference so that results are
// reproducible against its CPU render.  This translation unit set is compiled with -fmad=false (nvcc) and
// -ffp-contract=off (host) so that a*b+c is two roundings, as in the reference's x86-64 build.
// Reference: /root/reference/include/common/vector.h:48-337 (Vec3/Point3/Normal ops, createCs__),
//            /root/reference/include/utility/util_math_optimizations.h:99-270 (fast math),
//            /root/reference/src/common/vector.cc:86-141 (refract__, fresnel__).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#define YG_HD __host__ __device__ __forceinline__
#define YG_D __device__ __forceinline__
#define YG_DN static __device__ __noinline__

namespace yg {

struct V3 { float x, y, z; };
struct Col { float r, g, b; };

YG_HD V3 mk(float x, float y, float z) { V3 v; v.x = x; v.y = y; v.z = z; return v; }
YG_HD V3 mk(float s) { return mk(s, s, s); }
YG_HD V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
YG_HD V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
YG_HD V3 operator-(V3 a) { return mk(-a.x, -a.y, -a.z); }
YG_HD V3 operator*(float f, V3 b) { return mk(f * b.x, f * b.y, f * b.z); }   // vector.h:166-174
YG_HD V3 operator*(V3 b, float f) { return mk(f * b.x, f * b.y, f * b.z); }
YG_HD V3 operator/(V3 b, float f) { return mk(b.x / f, b.y / f, b.z / f); }   // vector.h:181-184
YG_HD float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }    // vector.h:161-164 (left to right)
YG_HD V3 cross(V3 a, V3 b) { return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); } // vector.h:201-204
YG_HD float lengthSqr(V3 a) { return (a.x * a.x + a.y * a.y) + a.z * a.z; }
YG_HD float fSqrt(float a) { return sqrtf(a); }                                 // util_math_optimizations.h:167-168,203-210
YG_HD float length(V3 a) { return fSqrt(lengthSqr(a)); }

// 1.0 / x evaluated in double and rounded to float equals the float division (innocuous double rounding, 53 >= 2*24+2)
YG_HD float rcp(float x) { return 1.0f / x; }

// Vec3::normalize vector.h:274-285
YG_HD V3 normalize(V3 v)
{
	float len = lengthSqr(v);
	if(len != 0.f)
	{
		len = rcp(fSqrt(len));
		v.x *= len; v.y *= len; v.z *= len;
	}
	return v;
}
// Vec3::normLen vector.h:63-73
YG_HD float normLen(V3 &v)
{
	float vl = lengthSqr(v);
	if(vl != 0.f)
	{
		vl = fSqrt(vl);
		const float d = rcp(vl);
		v.x *= d; v.y *= d; v.z *= d;
	}
	return vl;
}
// Vec3::reflect vector.h:313-320
YG_HD V3 reflect(V3 v, V3 n)
{
	const float vn = 2.0f * dot(v, n);
	return mk(vn * n.x - v.x, vn * n.y - v.y, vn * n.z - v.z);
}
// reflectDir__ vector.h:322-327
YG_HD V3 reflectDir(V3 n, V3 v)
{
	const float vn = dot(v, n);
	if(vn < 0) return -v;
	return (2 * vn) * n - v;
}
// createCs__ vector.h:339-357
YG_HD void createCs(V3 n, V3 &u, V3 &v)
{
	if((n.x == 0) && (n.y == 0))
	{
		u = (n.z < 0) ? mk(-1, 0, 0) : mk(1, 0, 0);
		v = mk(0, 1, 0);
	}
	else
	{
		const float d = rcp(fSqrt(n.y * n.y + n.x * n.x));
		u = mk(n.y * d, -n.x * d, 0);
		v = cross(n, u);
	}
}

YG_HD Col mkc(float r, float g, float b) { Col c; c.r = r; c.g = g; c.b = b; return c; }
YG_HD Col mkc(float s) { return mkc(s, s, s); }
YG_HD Col operator+(Col a, Col b) { return mkc(a.r + b.r, a.g + b.g, a.b + b.b); }
YG_HD Col operator*(Col a, Col b) { return mkc(a.r * b.r, a.g * b.g, a.b * b.b); }
YG_HD Col operator*(Col a, float f) { return mkc(a.r * f, a.g * f, a.b * f); }
YG_HD Col operator*(float f, Col a) { return mkc(a.r * f, a.g * f, a.b * f); }
YG_HD Col operator/(Col a, float f) { return mkc(a.r / f, a.g / f, a.b / f); }
// Rgb::clampProportionalRgb color.h:412-447
YG_HD Col clampProportional(Col c, float max_value)
{
	if(max_value > 0.f)
	{
		const float max_rgb = fmaxf(c.r, fmaxf(c.g, c.b));
		const float k = max_value / max_rgb;
		if(max_rgb > max_value)
		{
			if(c.r >= max_rgb) { c.r = max_value; c.g *= k; c.b *= k; }
			else if(c.g >= max_rgb) { c.g = max_value; c.r *= k; c.b *= k; }
			else { c.b = max_value; c.r *= k; c.g *= k; }
		}
	}
	return c;
}
YG_HD bool isBlack(Col c) { return c.r == 0 && c.g == 0 && c.b == 0; }
YG_HD float maxc(Col c) { return fmaxf(c.r, fmaxf(c.g, c.b)); }

YG_HD int f2i(float f) { union { float f; int i; } u; u.f = f; return u.i; }
YG_HD float i2f(int i) { union { float f; int i; } u; u.i = i; return u.f; }

// ---- fast math, util_math_optimizations.h -------------------------------------------------------------------------

// POLYEXP :99 — all-float Horner form
YG_HD float polyExp(float x)
{
	return (x * (x * (x * (x * (x * 1.8775767e-3f + 8.9893397e-3f) + 5.5826318e-2f) + 2.4015361e-1f) + 6.9315308e-1f) + 9.9999994e-1f);
}
// POLYLOG :100 — the constant 2.5988452 has no f suffix, so from there on the Horner form runs in double
YG_HD float polyLog(float x)
{
	float inner = x * (x * -3.4436006e-2f + 3.1821337e-1f) + -1.2315303f;
	double d = (double)(x * inner) + 2.5988452;
	d = (double)x * d + (double)-3.3241990f;
	d = (double)x * d + (double)3.1157899f;
	return (float)d;
}
// fExp2__ :116-130
YG_HD float fExp2(float x)
{
	x = fminf(x, 129.00000f);
	x = fmaxf(x, -126.99999f);
	int ipart = (int)(x - 0.5f);
	float fpart = x - (float)ipart;
	float expipart = i2f((int)((unsigned)(ipart + 127) << 23));
	return expipart * polyExp(fpart);
}
// fLog2__ :132-142
YG_HD float fLog2(float x)
{
	int i = f2i(x);
	float e = (float)(((i & 0x7F800000) >> 23) - 127);
	float m = i2f((i & 0x7FFFFF) | f2i(1.0f));
	return polyLog(m) * (m - 1.0f) + e;
}
YG_HD float fPow(float a, float b) { return fExp2(fLog2(a) * b); }             // :176-183
YG_HD float fLog(float a) { return fLog2(a) * (float)0.69314718055994530942; } // :185-192
YG_HD float fExp(float a) { return fExp2((float)1.4426950408889634074 * a); }  // :194-201

// fSin__ :222-243 (comparisons against the double constants as written)
YG_HD float fSin(float x)
{
	const double M2PI = 6.28318530717958647692, MPI = 3.14159265358979323846;
	if((double)x > M2PI || (double)x < -M2PI) x -= ((int)(x * (float)0.15915494309189533577)) * (float)M2PI;
	if((double)x < -MPI) x += (float)M2PI;
	else if((double)x > MPI) x -= (float)M2PI;
	x = ((float)1.27323954473516268615 * x) - ((float)0.40528473456935108578 * x * fabsf(x));
	float result = 0.225f * (x * fabsf(x) - x) + x;
	if(result <= -1.0f) return -1.0f;
	else if(result >= 1.0f) return 1.0f;
	else return result;
}
YG_HD float fCos(float x) { return fSin(x + (float)1.57079632679489661923); }  // :245-252
// fAcos__/fAsin__ :255-270 (libm inside the domain)
YG_HD float fAcos(float x)
{
	if(x <= -1.0f) return (float)3.14159265358979323846;
	else if(x >= 1.0f) return 0.0f;
	else return acosf(x);
}
YG_HD float fAsin(float x)
{
	if(x <= -1.0f) return (float)-1.57079632679489661923;
	else if(x >= 1.0f) return (float)1.57079632679489661923;
	else return asinf(x);
}

// float * double-constant products that the reference writes as `s * M_2PI` etc. (evaluated in double, stored to float)
YG_HD float mulD(float a, double c) { return (float)((double)a * c); }
#define YG_M_PI 3.14159265358979323846
#define YG_M_2PI 6.28318530717958647692
#define YG_M_1_PI 0.31830988618379067154

// ---- refraction / Fresnel, src/common/vector.cc:86-141 -------------------------------------------------------------

YG_HD bool refract(V3 n, V3 wi, V3 &wo, float ior)
{
	V3 N = n, i = -wi;
	float eta = ior;
	float cos_v_n = dot(wi, n);
	if(cos_v_n < 0) { N = -n; cos_v_n = -cos_v_n; }
	else eta = rcp(ior);
	float k = 1 - eta * eta * (1 - cos_v_n * cos_v_n);
	if(k <= 0.f) return false;
	wo = eta * i + (eta * cos_v_n - fSqrt(k)) * N;
	wo = normalize(wo);
	return true;
}

YG_HD void fresnel(V3 i, V3 n, float ior, float &kr, float &kt)
{
	float eta = ior;
	V3 N = (dot(i, n) < 0) ? -n : n;
	float c = dot(i, N);
	float g = eta * eta + c * c - 1;
	if(g <= 0) g = 0;
	else g = fSqrt(g);
	float aux = c * (g + c);
	// kr = ((0.5 * (g-c)*(g-c)) / ((g+c)*(g+c))) * (1 + ((aux-1)*(aux-1)) / ((aux+1)*(aux+1)));  0.5 and 1 are double/int literals:
	double t0 = (0.5 * (double)(g - c)) * (double)(g - c);
	double t1 = t0 / (double)((g + c) * (g + c));
	float q = ((aux - 1) * (aux - 1)) / ((aux + 1) * (aux + 1));
	kr = (float)(t1 * (double)(1 + q));
	if(kr < 1.0f) kt = 1 - kr;
	else kt = 0;
}

// FACE_FORWARD material.h:33
YG_HD V3 faceForward(V3 ng, V3 n, V3 i) { return (dot(ng, i) < 0) ? -n : n; }

// addMod1__ util_sample.h:183-187
YG_HD float addMod1(float a, float b) { float s = a + b; return s > 1 ? s - 1.f : s; }

// sampleCosHemisphere__ util_sample.h:45-54
YG_HD V3 sampleCosHemisphere(V3 n, V3 ru, V3 rv, float s1, float s2)
{
	if(s1 >= 1.0f) return n;
	float z1 = s1;
	float z2 = mulD(s2, YG_M_2PI);
	return (ru * fCos(z2) + rv * fSin(z2)) * fSqrt(1.0f - z1) + n * fSqrt(z1);
}
// sampleCone__ util_sample.h:80-86
YG_HD V3 sampleCone(V3 d, V3 u, V3 v, float max_cos_ang, float s1, float s2)
{
	float cos_ang = 1.f - (1.f - max_cos_ang) * s2;
	float sin_ang = fSqrt(1.f - cos_ang * cos_ang);
	float t1 = mulD(s1, YG_M_2PI);
	return (u * fCos(t1) + v * fSin(t1)) * sin_ang + d * cos_ang;
}

} // namespace yg
