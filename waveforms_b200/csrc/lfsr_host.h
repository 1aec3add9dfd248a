// lfsr_host.h -- host-side GF(2) algebra of the Galois LFSR step (jump-ahead tables).
#pragma once
#include <stdint.h>

namespace wf {

struct Gf2Mat { uint64_t col[64]; };  // column i = image of unit vector e_i

uint64_t lfsr_mask(int degree);
void gf2_step_matrix(uint64_t mask, int degree, Gf2Mat& a);
uint64_t gf2_apply(const Gf2Mat& a, int degree, uint64_t v);
void gf2_mul(const Gf2Mat& a, const Gf2Mat& b, int degree, Gf2Mat& c);
void gf2_step_powers(uint64_t mask, int degree, Gf2Mat out[64]);   // A^(2^j)
void gf2_step_power(uint64_t mask, int degree, uint64_t steps, Gf2Mat& out);  // A^steps
uint64_t lfsr_jump(uint64_t mask, int degree, uint64_t state, uint64_t steps);

}  // namespace wf
