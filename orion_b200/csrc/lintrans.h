// lintrans.h -- hoisted baby-step / giant-step diagonal matrix-vector products.
// Replaces LG/lineartransform.go:26-211 (lintrans.NewTransformation / Encode / Evaluator.EvaluateNew in Lattigo).
#pragma once
#include "scheme.h"

namespace ob {

struct LinearTransform {
    int level = 0, log_slots = 0, N1 = 1;
    scale_t scale = 1;
    int W = 0;                                        // level+1+nP limbs per encoded diagonal
    std::vector<int> diag_idx;                        // non-zero diagonals (mod slots), sorted
    std::map<int, int> slot_of;                       // diagonal index -> position in `diags`
    std::map<int, std::vector<int>> index;            // giant step j -> baby steps i (sorted)
    std::vector<int> babies, giants;                  // sorted distinct rotations
    DevBuf diags;                                     // [nd][W][N] NTT domain over QP, Montgomery form
    std::map<int, DevBuf> pairs;                      // giant j -> int2[(diag slot, rot slot)]
    bool has_data() const { return diags.p != nullptr; }
};

// BSGS planning (Lattigo lintrans.FindBestBSGSRatio / BSGSIndex, recalled)
void bsgs_index(const std::vector<int>& diags, int slots, int N1, std::map<int, std::vector<int>>& index,
                std::vector<int>& giants, std::vector<int>& babies);
int find_best_bsgs_ratio(const std::vector<int>& diags, int slots, int log_max_ratio);

// diagonals: idx -> slots complex values (not pre-rotated).  Encoded at `scale` into QP at `level`.
std::unique_ptr<LinearTransform> lt_build(Scheme& s, const std::map<int, std::vector<double2>>& diagonals, int level,
                                          scale_t scale, int log_slots, int log_bsgs_ratio, bool encode = true);
// encode = false (io_mode "load", LG/lineartransform.go:76-84): the BSGS plan only, with empty diagonals that
// LoadPlaintextDiagonal fills in later; the map's values are ignored.
std::unique_ptr<Ciphertext> lt_apply(Scheme& s, const LinearTransform& lt, const Ciphertext& in);
std::vector<u64> lt_galois_elements(Scheme& s, const LinearTransform& lt);

// handle table behind the C ABI (ltHeap in LG/lineartransform.go:15)
int lt_push(std::unique_ptr<LinearTransform> lt);
LinearTransform& lt_get(int id);
void lt_delete(int id);
void lt_reset();

}  // namespace ob
