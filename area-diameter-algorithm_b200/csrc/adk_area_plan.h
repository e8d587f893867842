// Table-slot planning of the area path's row sweep (host side, no CUDA): which resident cluster gets how much of the DP
// workspace and how long every anchor's ring of level bands is.  Kept free of CUDA so that tests/ can compile and check it
// on a machine without a GPU (tests/dropin/slot_plan_check.cpp).
#pragma once
#include <algorithm>
#include <cstdint>
#include <vector>

namespace adk
{
struct AnchorShape { uint32_t a, U, cmax; };      // points above the anchor (rows), table entries (sources), largest c of a record

struct SlotPlan
{
    uint64_t slots = 0;                 // resident clusters that get a slot (the launch has slots x CS CTAs)
    std::vector<uint64_t> off;          // slots + 1 entries: slot k is [off[k], off[k + 1]) in table entries, multiples of 32
    std::vector<uint32_t> W;            // per anchor: levels in its ring (a multiple of `band`)
};

// Cluster k starts with anchor k (the anchors arrive heaviest first) and then pulls from the queue behind the first `slots`
// anchors, so its slot must hold anchor k and whatever the queue may still hand out: slot k = max(need of anchor k, largest need
// in the tail).  Sizing every slot for the single largest anchor (round 1) took 86-120 GB at n = 1000 for two CTAs per SM.
//   need  = the shortest ring the sweep can run with: ceil(cmax / band) + CS bands (reach back cmax levels from the oldest band
//           in flight), never more than the levels an anchor with this many rows can fill;
//   full  = every level the anchor can fill: a ring that never wraps leaves the table complete when the anchor ends, and
//           the cluster can then walk its polygon back on the spot (end of k_area_rows) instead of the winner being re-run
//           after the solve.  Slots grow from need towards full as far as the memory budget allows (HBM is idle otherwise).
// An anchor lengthens its ring to fill the slot it will run in: its own for the first `slots` anchors, the smallest slot for the
// others.  Returns false when not even one slot fits the budget.
inline bool plan_area_slots(const std::vector<AnchorShape>& anchors, uint64_t K, uint32_t CS, uint64_t resident, uint64_t budget_entries,
                            uint32_t band, SlotPlan& out)
{
    const uint32_t cnt = (uint32_t) anchors.size();
    out.slots = 0; out.off.assign(1, 0); out.W.assign(cnt, band);
    if (cnt == 0) return true;
    std::vector<uint64_t> need(cnt), full(cnt), nbmax(cnt);
    for (uint32_t t = 0; t < cnt; ++t)
    {
        const AnchorShape& A = anchors[t];
        const uint64_t top_level = std::min<uint64_t>(K, (uint64_t) A.a + 2);
        nbmax[t] = std::max<uint64_t>(1, (top_level >= 3 ? top_level - 2 + band - 1 : 1) / band);
        const uint64_t wb = std::min<uint64_t>(((uint64_t) A.cmax + band - 1) / band + CS, nbmax[t]);
        need[t] = wb * band * (uint64_t) A.U;
        full[t] = nbmax[t] * band * (uint64_t) A.U;
    }
    uint64_t slots = std::min<uint64_t>(cnt, std::max<uint64_t>(resident / std::max<uint32_t>(CS, 1u), 1));
    std::vector<uint64_t> slot_size;
    uint64_t total_need = 0;
    for (;; --slots)
    {
        uint64_t tail_need = 1;
        for (uint32_t t = (uint32_t) slots; t < cnt; ++t) tail_need = std::max(tail_need, need[t]);
        slot_size.assign(slots, 0);
        total_need = 0;
        for (uint32_t k = 0; k < slots; ++k) { slot_size[k] = (std::max(need[k], tail_need) + 31) & ~(uint64_t) 31; total_need += slot_size[k]; }
        if (total_need <= budget_entries || slots == 1) break;
    }
    if (total_need > budget_entries) return false;
    {
        uint64_t tail_full = 1;
        for (uint32_t t = (uint32_t) slots; t < cnt; ++t) tail_full = std::max(tail_full, full[t]);
        uint64_t total_want = 0;
        std::vector<uint64_t> want(slots);
        for (uint32_t k = 0; k < slots; ++k) { want[k] = std::max(slot_size[k], (std::max(full[k], tail_full) + 31) & ~(uint64_t) 31); total_want += want[k]; }
        const double f = total_want <= budget_entries ? 1.0 : (double) (budget_entries - total_need) / (double) (total_want - total_need);
        for (uint32_t k = 0; k < slots; ++k)
            slot_size[k] = std::min<uint64_t>(want[k], slot_size[k] + ((uint64_t) ((double) (want[k] - slot_size[k]) * f) & ~(uint64_t) 31));
    }
    out.slots = slots;
    out.off.assign(slots + 1, 0);
    uint64_t min_slot = ~0ull;
    for (uint32_t k = 0; k < slots; ++k) { out.off[k + 1] = out.off[k] + slot_size[k]; min_slot = std::min(min_slot, slot_size[k]); }
    for (uint32_t t = 0; t < cnt; ++t)
    {
        const AnchorShape& A = anchors[t];
        const uint64_t room = t < slots ? slot_size[t] : min_slot;
        const uint64_t wb = A.U ? std::min<uint64_t>(nbmax[t], room / ((uint64_t) A.U * band)) : 1;
        out.W[t] = (uint32_t) (std::max<uint64_t>(wb, 1) * band);
    }
    return true;
}
}  // namespace adk
