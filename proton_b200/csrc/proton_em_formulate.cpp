// proton_em_formulate.cpp — the discretiser of a whole grid as one host call.
//
// What Matrix_Formulation_Class.matrix_formulation does per line in step mode (matrix_formulation_func_cli.py:24-299)
// and PROTON_cli.analyze_line_item's choice of step (PROTON_cli.py:1000-1014), over a flat segment table:
//   dx = step * 1e-6;  per segment  length = len_um * 1e-6,  "not dense" if length < dx,  points = round(length / dx)
//   (Python's round: half to even = rint), nx_total += points;  a segment whose constant coordinate equals its
//   predecessor's is joined to it by a via: vias[nx_total - points] = segment number (a dict: a repeated key keeps the
//   last segment), nx_total -= 1;  B gets (0, 0) = +b', per via (key-1, seg-2) = -b', (key-1, seg-1) = +b', and
//   (N-1, S-1) = -b', later assignments overwriting earlier ones;  Acoef = kappa / dx^2, b' = kappa b / dx;
//   current densities = float32(current / width^2).
// Host arithmetic only (IEEE double, no contraction: the file is compiled with -ffp-contract=off), so the arrays are
// bit-identical to the NumPy formulation in proton_b200/formulation.py (formulate_table) and to the per-line twin
// formulate_segments; tests/test_formulation.py compares all three on the reference-generated goldens.
// Needs no context and no GPU.
#include "../../include/proton_em.h"

#include <cmath>
#include <cstdint>
#include <cstdlib>

namespace {

struct Via { int64_t key, seg; };

inline void put(int32_t* rows, int32_t* cols, double* vals, int64_t first, int64_t& n, int64_t i, int64_t j, double v) {
  // a later assignment to the same position replaces the earlier one; the list is sorted, so only the last entry
  // of the line can be at the same position
  if (n > first && rows[n - 1] == (int32_t)i && cols[n - 1] == (int32_t)j) { vals[n - 1] = v; return; }
  rows[n] = (int32_t)i; cols[n] = (int32_t)j; vals[n] = v; ++n;
}

}  // namespace

extern "C" int proton_em_formulate_segments(const proton_em_segments* t, const double* step_um, double kappa, double b,
                                            double width2, int32_t* num_nodes, double* a_coeff, int64_t* trip_off,
                                            int32_t* trip_row, int32_t* trip_col, double* trip_val, double* curden,
                                            int64_t* bad_line) {
  if (bad_line) *bad_line = -1;
  if (!t || !num_nodes || !a_coeff || !trip_off || !trip_row || !trip_col || !trip_val || !curden) return PROTON_EM_EINVAL;
  const int64_t L = t->num_lines;
  if (L <= 0) return PROTON_EM_FORM_CORRUPT;
  for (int64_t l = 0; l < L; ++l)
    if (t->seg_off[l + 1] <= t->seg_off[l]) { if (bad_line) *bad_line = l; return PROTON_EM_FORM_CORRUPT; }
  // pass 1: the step of every line and the density test over the whole table (the batch fails as a whole, like the
  // NumPy formulation, before anything is written)
  for (int64_t l = 0; l < L; ++l) {
    const int64_t k0 = t->seg_off[l], k1 = t->seg_off[l + 1];
    double step;
    if (step_um) step = step_um[l];
    else {
      double mn = t->length_um[k0];
      for (int64_t k = k0 + 1; k < k1; ++k) {                       // np.minimum: a NaN wins
        const double x = t->length_um[k];
        if (x < mn || std::isnan(x)) mn = x;
      }
      step = mn * 0.99;
    }
    const double dx = step * 1e-6;
    for (int64_t k = k0; k < k1; ++k)
      if (t->length_um[k] * 1e-6 < dx) { if (bad_line) *bad_line = l; return PROTON_EM_FORM_NOT_DENSE; }
    a_coeff[l] = dx;            // parked here until pass 2
  }
  int64_t n = 0;
  trip_off[0] = 0;
  for (int64_t l = 0; l < L; ++l) {
    const int64_t k0 = t->seg_off[l], k1 = t->seg_off[l + 1];
    const double dx = a_coeff[l];
    const double bcoef = kappa * b / dx;
    const int64_t first = n;
    int64_t nx = 0;
    Via pend{-1, 0};
    bool have = false;
    put(trip_row, trip_col, trip_val, first, n, 0, 0, bcoef);
    for (int64_t k = k0; k < k1; ++k) {
      const double length = t->length_um[k] * 1e-6;
      curden[k] = (double)(float)(t->current[k] / width2);
      const double q = length / dx;
      if (!(std::fabs(q) < 9.0e18)) { if (bad_line) *bad_line = l; return PROTON_EM_FORM_CORRUPT; }   // dx = 0, NaN
      const int64_t d = (int64_t)std::nearbyint(q);                  // Python's round(): half to even
      nx += d;
      if (k > k0 && t->named[k - 1]) {
        const bool same2 = t->lcoord[2 * k + 1] == t->rcoord[2 * k + 1];
        const int64_t cur = same2 ? t->lcoord[2 * k + 1] : t->lcoord[2 * k];
        const int64_t prev = same2 ? t->lcoord[2 * (k - 1) + 1] : t->lcoord[2 * (k - 1)];
        if (prev == cur) {
          const int64_t key = nx - d, seg = k - k0 + 1;
          if (have && pend.key != key) {                 // the run of equal keys ended: its last segment is the via
            if (pend.key < 1 || pend.seg < 2) { if (bad_line) *bad_line = l; return PROTON_EM_FORM_INDEX; }
            put(trip_row, trip_col, trip_val, first, n, pend.key - 1, pend.seg - 2, -bcoef);
            put(trip_row, trip_col, trip_val, first, n, pend.key - 1, pend.seg - 1, bcoef);
          }
          pend = Via{key, seg};
          have = true;
          nx -= 1;
        }
      }
    }
    if (have) {
      if (pend.key < 1 || pend.seg < 2) { if (bad_line) *bad_line = l; return PROTON_EM_FORM_INDEX; }
      put(trip_row, trip_col, trip_val, first, n, pend.key - 1, pend.seg - 2, -bcoef);
      put(trip_row, trip_col, trip_val, first, n, pend.key - 1, pend.seg - 1, bcoef);
    }
    if (nx <= 0 || nx > 0x7fffffffLL) { if (bad_line) *bad_line = l; return PROTON_EM_FORM_CORRUPT; }
    put(trip_row, trip_col, trip_val, first, n, nx - 1, (k1 - k0) - 1, -bcoef);
    // (a via entry may not lie outside the matrix: key - 1 <= N - 1 and seg - 1 <= S - 1 hold by construction)
    num_nodes[l] = (int32_t)nx;
    a_coeff[l] = kappa / (dx * dx);
    trip_off[l + 1] = n;
  }
  return PROTON_EM_OK;
}
