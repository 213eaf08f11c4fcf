// QMC / RNG helpers restated from the reference, bit for bit (integers) and operation for operation (doubles).
// Reference: /root/reference/include/utility/util_mcqmc.h:28-190 (Halton, riVdC__, riS__, riLp__, fnv32ABuf__, Random),
//            /root/reference/include/common/scr_halton.h:37-76 (scrHalton__ + prims__/inv_prims__),
//            /root/reference/src/common/faure_tables.cc (Faure permutations; regenerated here by Faure's recursive
//            construction in yg_build_faure_tables() and checked against the reference's tables by the oracle KATs).
#pragma once
#include "dev_math.cuh"

namespace yg {

#define YG_MULT_RATIO 0.00000000023283064365386962890625 /* 2^-32, util_mcqmc.h:91 */

YG_HD float clamp01f(float v) { return fmaxf(0.f, fminf(1.f, v)); }

// riVdC__ util_mcqmc.h:93-101
YG_HD float riVdC(uint32_t bits, uint32_t r = 0)
{
	bits = (bits << 16) | (bits >> 16);
	bits = ((bits & 0x00ff00ffu) << 8) | ((bits & 0xff00ff00u) >> 8);
	bits = ((bits & 0x0f0f0f0fu) << 4) | ((bits & 0xf0f0f0f0u) >> 4);
	bits = ((bits & 0x33333333u) << 2) | ((bits & 0xccccccccu) >> 2);
	bits = ((bits & 0x55555555u) << 1) | ((bits & 0xaaaaaaaau) >> 1);
	return clamp01f((float)((double)(bits ^ r) * YG_MULT_RATIO));
}
// riS__ util_mcqmc.h:103-108 (`1 << 31` is 0x80000000u)
YG_HD float riS(uint32_t i, uint32_t r = 0)
{
	for(uint32_t v = 0x80000000u; i; i >>= 1, v ^= v >> 1)
		if(i & 1) r ^= v;
	return clamp01f((float)((double)r * YG_MULT_RATIO));
}
// riLp__ util_mcqmc.h:110-115
YG_HD float riLp(uint32_t i, uint32_t r = 0)
{
	for(uint32_t v = 0x80000000u; i; i >>= 1, v |= v >> 1)
		if(i & 1) r ^= v;
	return clamp01f((float)((double)r * YG_MULT_RATIO));
}
// fnv32ABuf__ util_mcqmc.h:147-162 (bytes of the value, little endian)
YG_HD uint32_t fnv32a(uint32_t value)
{
	uint32_t hash = 0x811c9dc5u;
	for(int i = 0; i < 4; i++)
	{
		hash ^= (value >> (8 * i)) & 0xffu;
		hash *= 0x01000193u;
	}
	return hash;
}

// Halton util_mcqmc.h:28-85 — incremental generator, double precision.  BASE is a compile-time constant so that `% BASE`
// and `/ BASE` become multiply-shift sequences; the arithmetic (order of the double additions and multiplications, the
// 0.9999999999 guard) is the reference's.
template<uint32_t BASE>
struct HaltonB
{
	double value;
	YG_HD void setStart(uint32_t i)
	{
		const double inv_base = 1.0 / (double)BASE;
		double factor = inv_base;
		value = 0.0;
		while(i > 0)
		{
			const uint32_t d = i % BASE;
			// (double)d * factor without the int -> double conversion (a slow instruction inside a 20-step dependent chain): 2 * factor is exact, so
			// factor + factor == 2.0 * factor, and value + 0.0 == value; the sum below has the bits of the reference's `value += (double)d * factor`
			if(BASE == 3) value += (d == 0 ? 0.0 : (d == 1 ? factor : factor + factor));
			else value += (double)d * factor;
			i /= BASE;
			factor *= inv_base;
		}
	}
	YG_HD float getNext()
	{
		const double inv_base = 1.0 / (double)BASE;
		double r = 0.9999999999 - value;
		if(inv_base < r) value += inv_base;
		else
		{
			double hh = 0.0, h = inv_base;
			while(h >= r) { hh = h; h *= inv_base; }
			value += hh + h - 1.0;
		}
		return fmaxf(0.f, fminf(1.f, (float)value));
	}
};

// Base 2.  Every intermediate of the reference's loops is exact in double here: the digits are bits, `factor` is a power of
// two, and a partial sum has at most 33 significant bits (32-bit start index plus the increments of one light/BSDF loop).
// The 1e-10 guard of getNext cannot change a branch either, because 1 - value and the candidate power of two differ by a
// multiple of 2^-33 = 1.16e-10.  Element i of the sequence is therefore exactly bitreverse64(i) * 2^-64, which is what this
// computes — same bits as the incremental generator (the exact-render tests pin it), without the loops.
struct Halton2
{
	unsigned long long idx;
	YG_HD void setStart(uint32_t i) { idx = i; }
	__device__ __forceinline__ float getNext()
	{
		++idx;
		const double value = (double)__brevll(idx) * 5.42101086242752217e-20; // 2^-64
		return fmaxf(0.f, fminf(1.f, (float)value));
	}
};

// Random (multiply-with-carry) util_mcqmc.h:173-190
struct Mwc
{
	uint32_t x, c;
	YG_HD void seed(uint32_t s) { x = 30903u; c = s; }
	YG_HD double next()
	{
		const uint32_t ya = 1791398085u, yah = ya >> 16, yal = ya & 65535u;
		const uint32_t xh = x >> 16, xl = x & 65535u;
		x = x * ya + c;
		c = xh * yah + ((xh * yal) >> 16) + ((xl * yah) >> 16);
		if(xl * yal >= ~c + 1) c++;
		return (double)x * YG_MULT_RATIO;
	}
};

// ---- scrambled Halton --------------------------------------------------------------------------------------------
#define YG_SCR_DIMS 50
#define YG_FAURE_TOTAL 4896 /* >= 4886 = 3 + sum of primes 5..227: dims 0..2 share the base-3 table (faure_tables.cc:437) */

struct FaureTables
{
	int offset[YG_SCR_DIMS];      // start of sigma for dimension d
	int base[YG_SCR_DIMS];        // prims__[d]
	double inv_base[YG_SCR_DIMS]; // inv_prims__[d] (9-decimal literals of the reference, NOT 1.0/base)
	unsigned long long magic[YG_SCR_DIMS]; // floor(2^64 / base) + 1: floor(n / base) == mulhi64(n, magic) for every 32-bit n (base >= 2)
	short sigma[YG_FAURE_TOTAL];
	double sigma_d[YG_FAURE_TOTAL];        // the same permutation values as doubles, so that the digit loop loads its addend instead of converting it
};

// scrHalton__ scr_halton.h:52-76.  `tab` lives in global memory on the device (indexed per thread: L1-cached loads, not the constant bank).
YG_HD double scrHaltonT(const FaureTables &tab, int dim, uint32_t n)
{
	double value = 0.0;
	const double *sigma = tab.sigma_d + tab.offset[dim];
	const uint32_t base = (uint32_t)tab.base[dim];
	// n % base by an exact integer reciprocal: with M = floor(2^64 / base) + 1 the error of n * M / 2^64 against n / base is below n / 2^64 < 2^-32, and the
	// fractional part of n / base is at most 1 - 1/base, so the floor is that of n / base for every 32-bit n.  The reference's own digit extraction is NOT an
	// integer division (dn *= f with the 9-decimal f, n = (unsigned)dn) and is kept as it is; only `n % base` and the int -> double conversion of the permuted
	// digit are replaced by bit-identical cheaper forms (4 conversions per digit before, 1 now).
	const unsigned long long magic = tab.magic[dim];
	double f, factor, dn = (double)n;
	f = factor = tab.inv_base[dim];
	while(n > 0)
	{
#ifdef __CUDA_ARCH__
		const uint32_t q = (uint32_t)__umul64hi((unsigned long long)n, magic);
#else
		const uint32_t q = (uint32_t)(((unsigned __int128)n * magic) >> 64);
#endif
		value += sigma[n - q * base] * factor;
		dn *= f;
		n = (uint32_t)dn;
		factor *= f;
	}
	return fmax(1.0e-36, fmin(1.0, value));
}

} // namespace yg

// Host: fills a FaureTables (csrc/faure.cpp)
void yg_build_faure_tables(yg::FaureTables &t);
