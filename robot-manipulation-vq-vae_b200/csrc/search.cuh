// Internal plan/workspace layout shared by the two nearest-code contractions (search.cu, search_tc.cu).
#pragma once
#include "common.cuh"

namespace lipvq {

// Header at the start of the search workspace (zeroed by a memset node every call).
struct SearchHdr {
  unsigned emax2_bits;  // max_k ||e_k||^2 as float bits (atomicMax on non-negative floats)
  int flagged;          // rows whose top-2 gap is inside the error bound -> exact rescore
  float mu_norm;        // ||mu||, mu = the centre subtracted from z and the codes before the contraction
  int pad[13];
};

struct SearchPlan {
  int algo;             // resolved LIPVQ_ALGO_*
  int ksplit;           // code range is cut in ksplit slices (grid.y) so small N still fills the SMs
  int codes_per_split;  // multiple of the code tile
  int Dp;               // tcgen05 path: D padded to a multiple of 64 (one 128-byte swizzle atom of bf16)
  int64_t Np;           // tcgen05 path: N padded to a multiple of 128
  int Kp;               // tcgen05 path: K padded to a multiple of the code tile
  size_t off_mu;        // centre mu[D] (fp32)
  size_t off_hdr, off_enorm, off_pbest, off_psecond, off_pidx, off_work;
  size_t off_zhi, off_zlo, off_ehi, off_elo;  // tcgen05 path: bf16 hi/lo splits
  size_t total;
};

SearchPlan make_search_plan(int64_t N, int D, int K, int metric, int algo);

// tcgen05 contraction (search_tc.cu): fills pbest/psecond/pidx [ksplit][N] exactly like the SIMT kernel.
// Returns LIPVQ_ERR_UNSUPPORTED when the shape is outside what the tensor path handles.
bool search_tc_supported(int64_t N, int D, int K);
int search_tc_launch(const float* z, int64_t N, int D, const float* cb, int K, const float* mu, const SearchPlan& plan,
                     char* ws, cudaStream_t stream);
// rigorous bound factor of the tcgen05 split-bf16 dot-product error (see search_tc.cu)
float search_tc_gamma(int D);
// resident-slice geometry of the tensor path for a shape: codes per slice, number of slices, padded latent dim
void search_tc_plan(int64_t N, int D, int K, int* codes_res, int* S, int* Dp);

}  // namespace lipvq
