// tests/emu/warp_emu.h -- TEST-ONLY lock-step emulation of one 32-lane warp on the host, for kernels whose lanes
// talk through shuffles and votes (bwa-mem2-arm_b200/csrc/bsw_wide.cuh).  Every lane is a ucontext coroutine; a
// collective publishes the lane's value, yields to the scheduler, and -- once all 32 lanes have arrived -- reads
// its partner's value.  Two value buffers alternate between consecutive collectives so that a lane that has already
// moved on cannot overwrite what a slower lane still has to read.
#pragma once
#include <ucontext.h>
#include <cstdint>
#include <cstdlib>
#include <functional>
#include <vector>

namespace warp_emu {

struct Warp {
    static constexpr int N = 32;
    ucontext_t sched, ctx[N];
    std::vector<char> stack[N];
    bool done[N];
    uint32_t val[2][N];
    int phase[N];
    int cur = -1;
    std::function<void(int)> body;
};
static thread_local Warp *g_w = nullptr;

static void trampoline()
{
    Warp *w = g_w;
    const int lane = w->cur;
    w->body(lane);
    w->done[lane] = true;
    swapcontext(&w->ctx[lane], &w->sched);
}

// runs body(lane) for lanes 0..31 in lock step
static inline void run(const std::function<void(int)> &body)
{
    Warp w;
    w.body = body;
    g_w = &w;
    for (int l = 0; l < Warp::N; ++l) {
        w.done[l] = false; w.phase[l] = 0;
        w.stack[l].resize(512 * 1024);
        getcontext(&w.ctx[l]);
        w.ctx[l].uc_stack.ss_sp = w.stack[l].data();
        w.ctx[l].uc_stack.ss_size = w.stack[l].size();
        w.ctx[l].uc_link = &w.sched;
        makecontext(&w.ctx[l], trampoline, 0);
    }
    for (;;) {                              // round robin: each lane runs until its next collective (or its end)
        bool any = false;
        for (int l = 0; l < Warp::N; ++l)
            if (!w.done[l]) { any = true; w.cur = l; swapcontext(&w.sched, &w.ctx[l]); }
        if (!any) break;
    }
    g_w = nullptr;
}

// publish, wait for the whole warp, return the buffer index the values of this collective live in
static inline int exchange(uint32_t v)
{
    Warp *w = g_w;
    const int lane = w->cur, b = w->phase[lane] & 1;
    w->val[b][lane] = v;
    w->phase[lane]++;
    swapcontext(&w->ctx[lane], &w->sched);          // every live lane reaches the same collective before anyone resumes
    w->cur = lane;
    return b;
}
static inline int lane_id() { return g_w->cur; }

} // namespace warp_emu

namespace bswk {
// the hooks bsw_wide.cuh declares for the host build; G = width of the shuffle segment
inline uint32_t w_shfl_up(uint32_t v, int d, int G)
{
    const int lane = warp_emu::lane_id(), b = warp_emu::exchange(v), src = lane - d;
    return (lane % G) - d >= 0 ? warp_emu::g_w->val[b][src] : v;
}
inline uint32_t w_shfl_xor(uint32_t v, int m, int G)
{
    const int lane = warp_emu::lane_id(), b = warp_emu::exchange(v);
    (void)G;
    return warp_emu::g_w->val[b][lane ^ m];
}
inline uint32_t w_shfl(uint32_t v, int src, int G)
{
    const int lane = warp_emu::lane_id(), b = warp_emu::exchange(v);
    return warp_emu::g_w->val[b][(lane / G) * G + (src % G)];
}
inline uint32_t w_redmax(uint32_t v, int G)
{
    const int lane = warp_emu::lane_id(), b = warp_emu::exchange(v), g0 = (lane / G) * G;
    uint32_t r = 0;
    for (int l = g0; l < g0 + G; ++l) r = warp_emu::g_w->val[b][l] > r ? warp_emu::g_w->val[b][l] : r;
    return r;
}
inline bool w_any(bool p)
{
    const int b = warp_emu::exchange(p ? 1u : 0u);
    bool r = false;
    for (int l = 0; l < warp_emu::Warp::N; ++l) r |= warp_emu::g_w->val[b][l] != 0;
    return r;
}
} // namespace bswk
