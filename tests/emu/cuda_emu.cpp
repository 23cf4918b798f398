/* tests/emu/cuda_emu.cpp — scheduler of the CPU SIMT emulator (see cuda_emu.h). TEST INFRASTRUCTURE ONLY. */
#include "cuda_emu.h"

extern "C" void emu_switch(void **from_sp, void *to_sp);
asm(".text\n.globl emu_switch\n.type emu_switch,@function\nemu_switch:\n"
    " pushq %rbp\n pushq %rbx\n pushq %r12\n pushq %r13\n pushq %r14\n pushq %r15\n"
    " movq %rsp, (%rdi)\n movq %rsi, %rsp\n"
    " popq %r15\n popq %r14\n popq %r13\n popq %r12\n popq %rbx\n popq %rbp\n ret\n"
    ".size emu_switch, .-emu_switch\n");

namespace emu {
State &S() { static State s; return s; }
void die(const char *msg) { fprintf(stderr, "cuda_emu: %s (block %u thread %u)\n", msg, S().bid.x, S().cur ? S().cur->tid.x : 0u); abort(); }
void yield() { State &s = S(); emu_switch(&s.cur->sp, s.sched_sp); }
static void fiber_entry() {
    State &s = S();
    (*s.body)();
    s.cur->done = true; s.progress++;
    for (;;) yield();
}
void warp_barrier() {
    Warp &w = W();
    unsigned gen = w.gen;
    if (++w.count == 32) { w.count = 0; w.gen++; S().progress++; return; }
    while (w.gen == gen) yield();
}
void launch(dim3 grid, dim3 block, size_t smem, const std::function<void()> &body) {
    State &s = S();
    unsigned nt = block.x * block.y * block.z;
    if (nt % 32) die("launch: block size must be a multiple of 32 in the emulator");
    s.bdim = block; s.gdim = grid; s.body = &body;
    s.dyn_smem.assign(smem + 16, 0xAB);
    static std::vector<char *> stacks;
    while (stacks.size() < nt) stacks.push_back((char *)aligned_alloc(64, STACK));
    for (unsigned bz = 0; bz < grid.z; bz++) for (unsigned by = 0; by < grid.y; by++) for (unsigned bx = 0; bx < grid.x; bx++) {
        s.bid = uint3{bx, by, bz};
        s.fibers.assign(nt, Fiber());
        s.warps.assign(nt / 32, Warp());
        s.cta_count = 0; s.cta_gen = 0;
        for (unsigned t = 0; t < nt; t++) {
            Fiber &f = s.fibers[t];
            f.tid = uint3{t % block.x, (t / block.x) % block.y, t / (block.x * block.y)};
            f.lane = t & 31; f.warp = t >> 5; f.stack = stacks[t];
            uintptr_t top = ((uintptr_t)f.stack + STACK) & ~(uintptr_t)15;
            void **sp = (void **)top;
            *--sp = nullptr;                 /* fake return address of fiber_entry */
            *--sp = (void *)&fiber_entry;    /* popped by emu_switch's ret */
            for (int i = 0; i < 6; i++) *--sp = nullptr;
            f.sp = sp;
        }
        unsigned long idle_rounds = 0, last = s.progress;
        for (;;) {
            bool all_done = true;
            for (unsigned t = 0; t < nt; t++) {
                Fiber &f = s.fibers[t];
                if (f.done) continue;
                all_done = false; s.cur = &f;
                emu_switch(&s.sched_sp, f.sp);
            }
            if (all_done) break;
            if (s.progress == last) { if (++idle_rounds > 2000000) { s.cur = &s.fibers[0]; die("deadlock: no barrier completed for 2M scheduler rounds (divergent collective or lost flag?)"); } }
            else { idle_rounds = 0; last = s.progress; }
        }
        s.cur = nullptr;
    }
}
}
void __syncthreads() {
    emu::State &s = emu::S();
    unsigned gen = s.cta_gen;
    if (++s.cta_count == (int)s.fibers.size()) { s.cta_count = 0; s.cta_gen++; s.progress++; return; }
    while (s.cta_gen == gen) emu::yield();
}
