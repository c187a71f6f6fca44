// Sequence-length balancing for dynamic micro-batches — HOST code (no kernels): the scheduler that decides which
// sequences share a micro-batch before the rmpad token stream is built.
//
// Restates verl/utils/seqlen_balancing.py:26-190 (karmarkar_karp + get_seqlen_balanced_partitions) with the
// reference's exact tie-breaking, because the partition a sequence lands in decides the order of every per-sequence
// tensor downstream (restore_dynamic_batch inverts it) and the parity bar for index work is bit-exact:
//   * a Set orders by (sum, item count, item list as (idx, len) tuples)                    seqlen_balancing.py:42-47
//   * a State keeps its k sets in DESCENDING order (stable sort)                           :57,71
//   * the priority queue pops the state with the LARGEST spread, ties: largest first set   :77-83
//   * the queue is Python's heapq: its sift order decides which of two equal-priority states pops first, so the
//     binary-heap sifts below follow heapq's (append + sift towards the root; pop = move the last leaf to the root,
//     sink it to a leaf along the smaller children, then sift it back up).
// Merging two states adds set i of the popped-first state to set k-1-i of the other (largest with smallest).
#include <algorithm>
#include <memory>
#include <utility>
#include <vector>

#include "common.h"

namespace vpg {
namespace {

struct KSet {
    long long sum = 0;
    std::vector<std::pair<long long, long long>> items;  // (idx, len) in insertion order
};

// Set.__lt__
inline bool set_lt(const KSet& a, const KSet& b) {
    if (a.sum != b.sum) return a.sum < b.sum;
    if (a.items.size() != b.items.size()) return a.items.size() < b.items.size();
    return a.items < b.items;  // lexicographic over (idx, len) pairs == Python list-of-tuples comparison
}

struct KState {
    std::vector<KSet> sets;  // descending
    long long spread() const { return sets.front().sum - sets.back().sum; }
    void sort_desc() {
        // sorted(self.sets, reverse=True): stable, equal elements keep their order
        std::stable_sort(sets.begin(), sets.end(), [](const KSet& a, const KSet& b) { return set_lt(b, a); });
    }
    void merge(const KState& other) {
        const size_t k = sets.size();
        for (size_t i = 0; i < k; ++i) {
            const KSet& o = other.sets[k - 1 - i];
            sets[i].items.insert(sets[i].items.end(), o.items.begin(), o.items.end());
            sets[i].sum += o.sum;
        }
        sort_desc();
    }
};

// State.__lt__: "a pops before b"
inline bool state_lt(const KState& a, const KState& b) {
    const long long sa = a.spread(), sb = b.spread();
    if (sa != sb) return sa > sb;
    return set_lt(b.sets[0], a.sets[0]);  // self.sets[0] > other.sets[0]  (reflected __lt__)
}

using Heap = std::vector<std::unique_ptr<KState>>;

void sift_towards_root(Heap& h, size_t startpos, size_t pos) {  // heapq._siftdown
    std::unique_ptr<KState> item = std::move(h[pos]);
    while (pos > startpos) {
        const size_t parent = (pos - 1) >> 1;
        if (state_lt(*item, *h[parent])) {
            h[pos] = std::move(h[parent]);
            pos = parent;
            continue;
        }
        break;
    }
    h[pos] = std::move(item);
}

void sink_then_sift(Heap& h, size_t pos) {  // heapq._siftup
    const size_t end = h.size(), start = pos;
    std::unique_ptr<KState> item = std::move(h[pos]);
    size_t child = 2 * pos + 1;
    while (child < end) {
        const size_t right = child + 1;
        if (right < end && !state_lt(*h[child], *h[right])) child = right;
        h[pos] = std::move(h[child]);
        pos = child;
        child = 2 * pos + 1;
    }
    h[pos] = std::move(item);
    sift_towards_root(h, start, pos);
}

void heap_push(Heap& h, std::unique_ptr<KState> s) {
    h.push_back(std::move(s));
    sift_towards_root(h, 0, h.size() - 1);
}

std::unique_ptr<KState> heap_pop(Heap& h) {
    std::unique_ptr<KState> last = std::move(h.back());
    h.pop_back();
    if (h.empty()) return last;
    std::unique_ptr<KState> top = std::move(h[0]);
    h[0] = std::move(last);
    sink_then_sift(h, 0);
    return top;
}

}  // namespace
}  // namespace vpg

extern "C" int vpg_seqlen_balanced_partitions(const int64_t* seqlens, int64_t n, int64_t k, int equal_size,
                                              int64_t* part_indices, int64_t* part_offsets) {
    using namespace vpg;
    VPG_CHECK_ARG(seqlens && part_indices && part_offsets, "null pointer argument");
    VPG_CHECK_ARG(k > 0, "k_partitions must be positive");
    VPG_CHECK_ARG(n >= k, "number of items:[%lld] < k_partitions:[%lld]", static_cast<long long>(n),
                  static_cast<long long>(k));
    VPG_CHECK_ARG(!equal_size || n % k == 0, "%lld %% %lld != 0", static_cast<long long>(n), static_cast<long long>(k));

    // sorted([(seqlen, i)])
    std::vector<std::pair<long long, long long>> order(static_cast<size_t>(n));
    for (int64_t i = 0; i < n; ++i) order[static_cast<size_t>(i)] = {seqlens[i], i};
    std::sort(order.begin(), order.end());

    Heap heap;
    heap.reserve(static_cast<size_t>(n));
    auto new_state = [&](size_t first, size_t count) {
        std::unique_ptr<KState> s(new KState);
        s->sets.resize(static_cast<size_t>(k));
        for (size_t i = 0; i < count; ++i) {
            const auto& [len, idx] = order[first + i];
            s->sets[i].items.emplace_back(idx, len);
            s->sets[i].sum += len;
        }
        s->sort_desc();
        return s;
    };
    if (equal_size) {
        for (size_t off = 0; off < order.size(); off += static_cast<size_t>(k))
            heap_push(heap, new_state(off, static_cast<size_t>(k)));
    } else {
        for (size_t i = 0; i < order.size(); ++i) heap_push(heap, new_state(i, 1));
    }
    while (heap.size() > 1) {
        std::unique_ptr<KState> s0 = heap_pop(heap);
        std::unique_ptr<KState> s1 = heap_pop(heap);
        s0->merge(*s1);
        heap_push(heap, std::move(s0));
    }
    const KState& fin = *heap[0];
    // _check_and_sort_partitions (:177-187): every partition non-empty, indices sorted inside a partition
    int64_t pos = 0;
    for (int64_t p = 0; p < k; ++p) {
        const KSet& s = fin.sets[static_cast<size_t>(p)];
        if (s.items.empty()) return set_error(VPG_ERR_INVALID, "the %lld-th partition is empty", static_cast<long long>(p));
        if (equal_size && static_cast<int64_t>(s.items.size()) * k != n)
            return set_error(VPG_ERR_INVALID, "%lld * %lld != %lld", static_cast<long long>(s.items.size()),
                             static_cast<long long>(k), static_cast<long long>(n));
        part_offsets[p] = pos;
        for (const auto& it : s.items) part_indices[pos++] = it.first;
        std::sort(part_indices + part_offsets[p], part_indices + pos);
    }
    part_offsets[k] = pos;
    return VPG_OK;
}
