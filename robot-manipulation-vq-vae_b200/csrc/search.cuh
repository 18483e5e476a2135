// Internal plan/workspace layout shared by the two nearest-code contractions (search.cu, search_tc.cu).
#pragma once
#include "common.cuh"

namespace lipvq {

// Header at the start of the search workspace (zeroed by a memset node every call).
struct SearchHdr {
  unsigned emax2_bits;  // max_k ||e_k||^2 as float bits (atomicMax on non-negative floats)
  int flagged;          // rows whose top-2 gap is inside the error bound -> exact rescore
  float mu_norm;        // ||mu||, mu = the centre subtracted from z and the codes before the contraction
  unsigned excl_inv_bits;  // fused path: max over codes excluded from the fp16 contraction of 1/||e_k-mu|| (0 = none)
  int flagged2;         // fused path: rows the 3-term pass could not certify either -> exact rescore
  int arrived;          // single-launch first pass: CTAs that have published their slice of the prepared codebook
  int done;             // single-launch first pass with its own exact scan: CTAs that have finished (statistics ticket)
  int ncand, nall;      // ... its statistics: codes evaluated exactly / rows whose every code was a candidate
  int pad[7];
};

struct SearchPlan {
  int algo;             // resolved LIPVQ_ALGO_*
  int ksplit;           // code range is cut in ksplit slices (grid.y) so small N still fills the SMs
  int codes_per_split;  // multiple of the code tile
  int Dp;               // tcgen05 path: D padded to a multiple of 64 (one 128-byte swizzle atom of fp16)
  int64_t Np;           // N padded to a multiple of 128
  int Kp;               // tcgen05 path: K padded to a multiple of the code tile
  size_t off_mu;        // centre mu[D] (fp32)
  size_t off_hdr, off_enorm, off_pbest, off_psecond, off_pidx, off_work;
  size_t off_work2;                  // tcgen05 path: second-level queue
  size_t off_ehi, off_elo;           // tcgen05 path: fp16 hi/lo of -2(e-mu), [Kp][Dp]
  size_t off_img;                    // single-launch first pass: image of its shared-memory code tiles + norm slab
  size_t off_prep;                   // prepared codebook state (search_prep_kernel); its position does not depend on N
  size_t total;
};

SearchPlan make_search_plan(int64_t N, int D, int K, int metric, int algo);

// Fused tcgen05 search (search_fused.cu): codebook prep, fp16 pass over all rows, 3-term pass over the queued rows.
// Writes idx_out for every certified row; *final_queue / *final_count name the rows left for the exact scan.
bool search_fused_supported(int64_t N, int D, int K);
void search_fused_plan(int D, int K, int* Kp, int* Dp);
// (*final_count == nullptr: every row has been settled, nothing is left for the exact scan)
int search_fused_launch(const float* z, int64_t N, int D, const float* cb, int K, int metric, const SearchPlan& plan,
                        char* ws, int64_t* idx_out, cudaStream_t stream, const int** final_queue,
                        const int** final_count, int32_t* stats_out);
// Single-launch first pass for codebooks that fit shared memory (search_p1.cu): codebook preparation + fp16 tcgen05 pass
// with the packed-key fold.  Uncertified rows are appended to queue (*queue_count must be zero on entry); mu / enorm /
// ehi / elo ([Kp_pub][Dp] fp16) / hdr are published for the later passes.
bool search_p1_supported(int64_t N, int D, int K);
bool search_p1_fits(int64_t N, int D, int K);  // geometry only (workspace planning)
size_t search_p1_image_bytes(int D, int K);
// prepared_state: nullptr (the kernel prepares the codebook itself) or the state search_prep_launch wrote for this codebook.
// tail: the kernel settles its own uncertified rows (exact scan); queue then needs search_p1_queue_rows(N) ints.
int search_p1_launch(const float* z, int64_t N, int D, const float* cb, int K, int metric, int Kp_pub, float* mu,
                     float* enorm, void* ehi, void* elo, void* img, SearchHdr* hdr, int* queue, int* queue_count, int64_t* idx_out,
                     cudaStream_t stream, void* prepared_state, int tail, int32_t* stats_out, float* queue_thr);
size_t search_p1_state_bytes(int D, int K);
int64_t search_p1_queue_rows(int64_t N);
int search_prep_launch(const float* cb, int K, int D, int metric, const float* center, int64_t n_center, void* state,
                       cudaStream_t stream);
bool search_p1_tail();  // the single-launch pass settles its own rows (default; LIPVQ_P1_TAIL=0: later passes as in round 1)
int search_skip_rows(int64_t N, int D, int K);
int search_best_values(const float* z, int64_t N, int D, const float* cb, int metric, const int64_t* idx, float* best_out,
                       cudaStream_t stream);

}  // namespace lipvq
