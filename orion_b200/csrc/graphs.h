// graphs.h -- "one ciphertext in, one ciphertext out" evaluator sequences as captured CUDA graphs (used for Bootstrap).
//
// At logN = 13 the host enqueues a bootstrap's 1 036 launches and ~650 pool calls in 3.8 ms, and a ResNet-20 inference is 28 of
// those plus 138 polynomial evaluations (0.8 ms of host time each): the single host thread is busy 93 % of an inference.  Such a
// sequence is the same launch sequence whenever its key repeats (slot count / polynomial, input level, input scale), so after one
// eager call (which builds every lazily created table) it is captured and replayed: the input is copied into a persistent staging
// ciphertext, the graph runs -- kernels, side-stream forks, programmatic edges; its scratch comes from a CaptureArena (common.h),
// so it holds no memory nodes: cudaGraphLaunch costs 6 us of host time -- and the result is copied out of a persistent result
// buffer.  Anything that cannot be captured (a table built lazily after all, a synchronisation) ends the capture
// (capture_abort): the call is repeated eagerly and the capture is tried again, at most three times.
// One key owns a few INSTANCES: an instance is one executable graph with its own staging and scratch and runs one call at a time.
// A call takes the instance it used last on this stream (stream order keeps it safe), else one whose previous launch has
// completed, else a new one (up to 4: the ciphertexts of a tensor are processed concurrently on the dataflow lanes), else it queues
// behind the least recently used one.  Results are bit-identical with the eager path (tools/graph_check.py, tests).
// Measured: ResNet-20 at logN=13 0.285 -> 0.272 s/image with the 28 bootstraps replayed; EvaluatePolynomial was tried the
// same way (138 calls, ~100 keys) and removed: no further gain -- with the bootstraps' host time gone the replay is bound by the
// GPU (K images in lockstep: 3.9 images/s with or without graphs) -- for 58 GB of arenas.
#pragma once
#include <functional>
#include "scheme.h"

namespace ob {

struct UnaryGraph {
    int state = 1, tries = 0;                             // 1 to be captured, 2 ready, -1 uncapturable
    unsigned long long epoch = 0, used = 0, launches = 0;
    double alg = 0;
    cudaGraphExec_t exec = nullptr;
    cudaStream_t last_stream = nullptr;
    cudaEvent_t done = nullptr;                           // completion of the last launch (the result copy included)
    DevBuf in_stage, out_stage, arena;                    // persistent: input copy, result copy, scratch of the captured sequence
    size_t arena_cap = 0, held = 0;                       // `held`: its share of the graphs' memory budget (graphs.cu)
    ~UnaryGraph();
};
struct UnaryGraphSet {
    int seen = 0;
    size_t scratch = 0;                                   // pool high-water mark of the first (eager) call: sizes the arenas
    size_t out_bytes = 0; int out_level = 0, out_degree = 1, out_log_slots = 0; scale_t out_scale = 1;   // learnt from the first (eager) call
    std::vector<std::unique_ptr<UnaryGraph>> inst;
    unsigned long long clock = 0;
};
// `s` allocates the result; `eager(x)` is the sequence itself (run for real on first sight of the key, under capture afterwards)
std::unique_ptr<Ciphertext> run_unary_graphed(Scheme& s, UnaryGraphSet& set, const Ciphertext& in, size_t arena_cap, const char* what,
                                              const std::function<std::unique_ptr<Ciphertext>(const Ciphertext&)>& eager);

}  // namespace ob
