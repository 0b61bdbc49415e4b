// graphs.cu -- capture / replay of unary evaluator sequences (graphs.h).
#include "graphs.h"
#include <ctime>

namespace ob {

static size_t g_graph_bytes = 0;                             // staging + scratch held by all live instances
UnaryGraph::~UnaryGraph() { if (exec) cudaGraphExecDestroy(exec); if (done) cudaEventDestroy(done); g_graph_bytes -= held; }

std::unique_ptr<Ciphertext> run_unary_graphed(Scheme& s, UnaryGraphSet& set, const Ciphertext& in, size_t arena_cap, const char* what,
                                              const std::function<std::unique_ptr<Ciphertext>(const Ciphertext&)>& eager) {
    if (g_capturing) return eager(in);                              // nested inside another captured sequence: part of that graph
    static const bool dbg = std::getenv("ORION_B200_DEBUG") != nullptr;
    const cudaStream_t st = stream();
    if (set.seen++ == 0) {                                          // first sight of the key: eager; builds every lazy table
        device_bytes_mark();
        auto out = eager(in);
        set.scratch = device_bytes_since_mark();
        set.out_bytes = out->data.bytes; set.out_level = out->level; set.out_degree = out->degree;
        set.out_log_slots = out->log_slots; set.out_scale = out->scale;
        return out;
    }
    for (auto& gi : set.inst)                                       // shared state a graph points at was rebuilt: capture again
        if (gi->state == 2 && gi->epoch != g_graph_epoch) {
            OB_CUDA(cudaStreamSynchronize(gi->last_stream));
            if (gi->exec) { cudaGraphExecDestroy(gi->exec); gi->exec = nullptr; }
            gi->state = 1;
        }
    UnaryGraph* pick = nullptr;
    for (auto& gi : set.inst) if (gi->state > 0 && gi->last_stream == st) { pick = gi.get(); break; }
    if (!pick) for (auto& gi : set.inst) if (gi->state > 0 && (!gi->done || cudaEventQuery(gi->done) == cudaSuccess)) { pick = gi.get(); break; }
    cudaGetLastError();                                             // cudaErrorNotReady from the queries is not an error
    // every instance keeps its scratch for good: new instances stop when the graphs hold a quarter of the device's memory
    // (a caller that walks through many distinct (level, scale) keys then simply runs the later ones eagerly)
    static const size_t graph_budget = [] { size_t fr = 0, tot = 0; cudaMemGetInfo(&fr, &tot); return tot / 4; }();
    const size_t want = (set.scratch ? std::max<size_t>((size_t)32 << 20, set.scratch * 5 / 2) : arena_cap) + in.data.bytes + set.out_bytes;
    if (!pick && set.inst.size() < 4 && g_graph_bytes + want <= graph_budget) {
        g_graph_bytes += want;
        set.inst.push_back(std::make_unique<UnaryGraph>());
        set.inst.back()->held = want;
        // the arena reuses a block only on the stream that freed it and wastes up to half a block per reuse: 2.5 x the pool's
        // high-water mark of the eager call fits in practice (it doubles when it does not); `arena_cap` is the fallback
        pick = set.inst.back().get();
        pick->arena_cap = set.scratch ? std::max<size_t>((size_t)32 << 20, set.scratch * 5 / 2) : arena_cap;
    }
    if (!pick) {                                                    // all busy elsewhere: queue behind the least recently used
        for (auto& gi : set.inst) if (gi->state > 0 && (!pick || gi->used < pick->used)) pick = gi.get();
        if (!pick) return eager(in);                                // every instance uncapturable
        if (pick->done) OB_CUDA(cudaStreamWaitEvent(st, pick->done, 0));
    }
    UnaryGraph& g = *pick;
    g.used = ++set.clock;
    const size_t in_bytes = (size_t)(in.degree + 1) * (in.level + 1) * s.N * 8;
    if (g.state == 1) {
        if (!g.in_stage.p) { g.in_stage.alloc(in_bytes); g.out_stage.alloc(set.out_bytes); }
        if (!g.arena.p) g.arena.alloc(g.arena_cap);
        Ciphertext view;                                            // the staging buffer seen as the input ciphertext (not owned)
        view.level = in.level; view.degree = in.degree; view.log_slots = in.log_slots; view.scale = in.scale;
        view.data.p = g.in_stage.p; view.data.bytes = 0;
        const unsigned long long l0 = g_launches, epoch0 = g_graph_epoch;
        const double a0 = g_alg_bytes;
        cudaGraph_t graph = nullptr;
        CaptureArena A; A.base = (char*)g.arena.p; A.cap = g.arena_cap;
        bool ok = capture_begin(st, &A);
        if (ok) {
            static const bool graph_pdl = !(std::getenv("ORION_B200_GRAPH_PDL") && std::atoi(std::getenv("ORION_B200_GRAPH_PDL")) == 0);
            if (!graph_pdl) ++g_pdl_suppress;
            std::unique_ptr<Ciphertext> out = eager(view);
            if (!graph_pdl) --g_pdl_suppress;
            ok = out->data.bytes == set.out_bytes && out->level == set.out_level && out->scale == set.out_scale;
            if (ok) OB_CUDA(cudaMemcpyAsync(g.out_stage.p, out->data.p, set.out_bytes, cudaMemcpyDeviceToDevice, st));
            out.reset();                                            // back to the arena: the graph holds no memory nodes
            graph = capture_end();
            ok = ok && graph != nullptr && epoch0 == g_graph_epoch && !A.overflow;
        }
        view.data.p = nullptr;
        if (ok) ok = cudaGraphInstantiate(&g.exec, graph, 0) == cudaSuccess;
        if (graph) cudaGraphDestroy(graph);
        if (!ok) {
            cudaGetLastError();
            g.exec = nullptr;
            g_launches = l0; g_alg_bytes = a0;
            OB_CUDA(cudaStreamSynchronize(st));                     // whatever ran eagerly after an aborted capture is done with the arena
            g.arena.release();
            if (A.overflow && g.arena_cap < ((size_t)16 << 30)) g.arena_cap *= 2;          // scratch did not fit: a larger arena next time
            else g.state = ++g.tries < 3 ? 1 : -1;                  // a table built lazily after all: it exists now, try again
            if (dbg) std::fprintf(stderr, "[graph] %s (level %d): capture %s\n", what, in.level,
                                  A.overflow ? "ran out of scratch, arena doubled" : "aborted, eager this time");
            return eager(in);
        }
        g.launches = g_launches - l0; g.alg = g_alg_bytes - a0;
        g_launches = l0; g_alg_bytes = a0;                          // nothing ran yet: the replay below counts
        g.epoch = g_graph_epoch; g.state = 2;
        if (dbg) std::fprintf(stderr, "[graph] %s (level %d): captured %llu kernel nodes, %zu MB of scratch\n", what, in.level, g.launches, A.top >> 20);
    }
    OB_CUDA(cudaMemcpyAsync(g.in_stage.p, in.data.p, in_bytes, cudaMemcpyDeviceToDevice, st));
    OB_CUDA(cudaGraphLaunch(g.exec, st));
    g_launches += g.launches; g_alg_bytes += g.alg;
    auto res = s.new_ct(set.out_level, set.out_degree, set.out_scale, set.out_log_slots);
    OB_CUDA(cudaMemcpyAsync(res->data.p, g.out_stage.p, set.out_bytes, cudaMemcpyDeviceToDevice, st));
    if (!g.done) OB_CUDA(cudaEventCreateWithFlags(&g.done, cudaEventDisableTiming));
    OB_CUDA(cudaEventRecord(g.done, st));
    g.last_stream = st;
    return res;
}

}  // namespace ob
