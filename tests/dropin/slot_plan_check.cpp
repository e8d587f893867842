// CPU check of the area path's table-slot planner (csrc/adk_area_plan.h): random anchor populations, every invariant the
// row sweep relies on.  Built and run by tests/test_boundary_cpu.py::test_area_slot_plan_invariants (g++, no CUDA).
#include "adk_area_plan.h"
#include <cstdio>
#include <random>

#define CHECK(cond) do { if (!(cond)) { std::fprintf(stderr, "case %d: %s failed (line %d)\n", cs, #cond, __LINE__); return 1; } } while (0)

int main()
{
    std::mt19937_64 rng(12345);
    const uint32_t band = 32;
    int planned = 0, refused = 0;
    for (int cs = 0; cs < 4000; ++cs)
    {
        const uint32_t cnt = 1 + (uint32_t) (rng() % (cs % 7 == 0 ? 1200 : 60));
        const uint32_t n = 3 + (uint32_t) (rng() % 4000);
        std::vector<adk::AnchorShape> A(cnt);
        for (uint32_t t = 0; t < cnt; ++t)
        {   // roughly heaviest first, like the anchors in y order, with noise; some empty anchors
            const uint32_t a = 2 + (uint32_t) ((uint64_t) (n - 2) * (cnt - t) / cnt);
            const uint32_t U = rng() % 11 == 0 ? 0u : (uint32_t) (rng() % ((uint64_t) a * a / 3 + 2));
            A[t] = adk::AnchorShape{a, U, (uint32_t) (rng() % (a + 1))};
        }
        const uint64_t K = rng() % 5 == 0 ? 3 + rng() % 40 : (uint64_t) n;
        const uint32_t CS = 1u << (rng() % 4);
        const uint64_t resident = 1 + rng() % 600;
        const uint64_t budget = (uint64_t) 1 << (18 + rng() % 22);       // 256 K .. 512 G entries
        adk::SlotPlan P;
        const bool ok = adk::plan_area_slots(A, K, CS, resident, budget, band, P);
        std::vector<uint64_t> need(cnt), nbmax(cnt);
        uint64_t max_need = 1;
        for (uint32_t t = 0; t < cnt; ++t)
        {
            const uint64_t top = std::min<uint64_t>(K, (uint64_t) A[t].a + 2);
            nbmax[t] = std::max<uint64_t>(1, (top >= 3 ? top - 2 + band - 1 : 1) / band);
            need[t] = std::min<uint64_t>((A[t].cmax + band - 1) / band + CS, nbmax[t]) * band * A[t].U;
            max_need = std::max(max_need, need[t]);
        }
        if (!ok)
        {   // a refusal is only allowed when the largest anchor alone does not fit
            CHECK(((max_need + 31) & ~31ull) > budget);
            ++refused;
            continue;
        }
        ++planned;
        const uint64_t S = P.slots;
        CHECK(S >= 1 && S <= cnt && S <= std::max<uint64_t>(resident / CS, 1));
        CHECK(P.off.size() == S + 1 && P.off[0] == 0 && P.W.size() == cnt);
        CHECK(P.off[S] <= budget);
        uint64_t tail_need = 0, min_slot = ~0ull;
        for (uint32_t t = (uint32_t) S; t < cnt; ++t) tail_need = std::max(tail_need, need[t]);
        for (uint64_t k = 0; k < S; ++k)
        {
            const uint64_t size = P.off[k + 1] - P.off[k];
            CHECK(P.off[k] % 32 == 0 && size % 32 == 0 && size >= 32);
            CHECK(size >= need[k] && size >= tail_need);                 // anchor k and anything the queue hands out later fit
            min_slot = std::min(min_slot, size);
        }
        for (uint32_t t = 0; t < cnt; ++t)
        {
            const uint64_t room = t < S ? P.off[t + 1] - P.off[t] : min_slot;
            CHECK(P.W[t] % band == 0 && P.W[t] >= band && P.W[t] / band <= nbmax[t]);
            CHECK((uint64_t) P.W[t] * A[t].U <= room);                   // the ring lies inside the slot
            CHECK((uint64_t) P.W[t] * A[t].U >= need[t]);                // and is at least the shortest ring the sweep can run with
        }
    }
    std::printf("slot plans: %d planned, %d refused\n", planned, refused);
    return planned > 1000 && refused > 0 ? 0 : 2;
}
