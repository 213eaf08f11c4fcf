// Faure permutations for scrHalton__ (reference: /root/reference/include/common/scr_halton.h:37-76 and the
// tables of /root/reference/src/common/faure_tables.cc).  The reference ships the permutations as literal tables;
// here they are regenerated with Faure's recursive construction
//   sigma_2 = (0 1);  b even: sigma_b = (2 sigma_{b/2}, 2 sigma_{b/2} + 1);
//   b odd, c = (b-1)/2: take sigma_{b-1}, add 1 to every value >= c, insert c in the middle
// and the oracle KAT "scrHalton" (tests/test_oracle_kat.py) pins them against the reference's own output.
#include "dev_qmc.cuh"
#include <vector>

static std::vector<int> faurePerm(int b)
{
	if(b == 2) return {0, 1};
	if(b % 2 == 0)
	{
		std::vector<int> h = faurePerm(b / 2), r;
		for(int v : h) r.push_back(2 * v);
		for(int v : h) r.push_back(2 * v + 1);
		return r;
	}
	const int c = (b - 1) / 2;
	std::vector<int> r = faurePerm(b - 1);
	for(int &v : r) if(v >= c) ++v;
	r.insert(r.begin() + c, c);
	return r;
}

void yg_build_faure_tables(yg::FaureTables &t)
{
	// prims__ / inv_prims__ exactly as written in scr_halton.h:37-50 (inverse rounded to 9 decimals)
	static const int prims[50] = {1, 2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37, 41, 43, 47, 53, 59, 61, 67,
	                              71, 73, 79, 83, 89, 97, 101, 103, 107, 109, 113, 127, 131, 137, 139, 149, 151, 157, 163, 167,
	                              173, 179, 181, 191, 193, 197, 199, 211, 223, 227};
	static const double inv_prims[50] = {1.000000000, 0.500000000, 0.333333333, 0.200000000, 0.142857143,
	                                     0.090909091, 0.076923077, 0.058823529, 0.052631579, 0.043478261,
	                                     0.034482759, 0.032258065, 0.027027027, 0.024390244, 0.023255814,
	                                     0.021276596, 0.018867925, 0.016949153, 0.016393443, 0.014925373,
	                                     0.014084507, 0.013698630, 0.012658228, 0.012048193, 0.011235955,
	                                     0.010309278, 0.009900990, 0.009708738, 0.009345794, 0.009174312,
	                                     0.008849558, 0.007874016, 0.007633588, 0.007299270, 0.007194245,
	                                     0.006711409, 0.006622517, 0.006369427, 0.006134969, 0.005988024,
	                                     0.005780347, 0.005586592, 0.005524862, 0.005235602, 0.005181347,
	                                     0.005076142, 0.005025126, 0.004739336, 0.004484305, 0.004405286};
	int pos = 0;
	for(int d = 0; d < YG_SCR_DIMS; ++d)
	{
		t.base[d] = prims[d];
		t.inv_base[d] = inv_prims[d];
		t.magic[d] = prims[d] >= 2 ? ~0ull / (unsigned long long)prims[d] + 1ull : 0ull;   // floor((2^64 - 1) / b) + 1 = floor(2^64 / b) + 1 for every odd b; 2^63 for b = 2, which divides exactly as well
		if(d < 2) { t.offset[d] = 0; continue; } // faure__[0], faure__[1] alias the base-3 table
		std::vector<int> p = faurePerm(prims[d]);
		t.offset[d] = pos;
		for(int v : p) t.sigma[pos++] = (short)v;
	}
	for(int k = 0; k < YG_FAURE_TOTAL; ++k) t.sigma_d[k] = (double)t.sigma[k];
}
