// TEST INFRASTRUCTURE ONLY: compiles the per-candidate math of compliancedex_b200/csrc/cdx_pg_math.cuh for the HOST
// (g++), so that the CPU test-suite can check the analytic FK Jacobian / Kabsch adjoint / closure gradient against
// the oracle's autograd without a GPU.  Nothing in the product loads this library; the product runs the same
// functions only inside the CUDA kernels of cdx_pregrasp.cu.
#include "../../compliancedex_b200/csrc/cdx_pg_math.cuh"
#include "../../compliancedex_b200/csrc/cdx_noise.cuh"

extern "C" {
int hc_build(int n_bodies, const int* parent, const int* dof, const int* axis, const float* sign, const float* rot,
             const float* xyz, int n_dofs, const int* tip_body, const float* tip_offset, cdx_hand_dev* out) {
  return cdx_build_chains(n_bodies, parent, dof, axis, sign, rot, xyz, n_dofs, tip_body, tip_offset, out);
}
int hc_hand_size() { return (int)sizeof(cdx_hand_dev); }
int hc_setup_size() { return (int)sizeof(cdx_pg_setup); }
// q [16] f64 -> tl [4,3] f64, J [4,3,CDX_MAX_CHAIN_DOF] f32
void hc_fk(const cdx_hand_dev* hand, const double* q, double* tl, float* J) {
  float q32[16];
  for (int i = 0; i < 16; i++) q32[i] = (float)q[i];
  for (int f = 0; f < 4; f++) {
    float tip[3];
    cdx_fk_finger(hand->finger[f], q32, tip, reinterpret_cast<float (*)[CDX_MAX_CHAIN_DOF]>(J + f * 3 * CDX_MAX_CHAIN_DOF));
    for (int r = 0; r < 3; r++) tl[f * 3 + r] = (double)tip[r];
  }
}
void hc_svd3(const double* A, double* U, double* S, double* V) { cdx_svd3(A, U, S, V); }
void hc_closure(const cdx_pg_setup* P, const cdx_hand_dev* hand, const double* q, const double* k, const double* T,
                const double* palm, const double* dist, const double* sd, const double* gmu, const double* gsd,
                const double* mean9, const double* gmean9, uint64_t seed, int it, int64_t b, double* total,
                double* tmargin, double* gq, double* gk, double* gT, double* gpalm, const double* tl_override) {
  double tl[4][3];
  float J[4][3][CDX_MAX_CHAIN_DOF];
  hc_fk(hand, q, &tl[0][0], &J[0][0][0]);
  if (tl_override)   // isolate the objective's derivation from float32 FK rounding
    for (int e = 0; e < 12; e++) (&tl[0][0])[e] = tl_override[e];
  cdx_pg_gpis G = {dist, sd, gmu, gsd, mean9, gmean9};
  auto nf = [=](int p, int e) { return cdx_kabsch_noise(seed, it, p, b, e); };
  cdx_pg_closure(*P, *hand, q, k, reinterpret_cast<const double (*)[3]>(T), palm, tl, J, G, nf, *total, tmargin, gq,
                 gk, reinterpret_cast<double (*)[3]>(gT), gpalm);
}
}
