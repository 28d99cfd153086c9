// Host work queues of the compute pass, re-expressed in C++.
// Mirrors src/utils.rs: QueueMap (:58-149), QueueSet (:151-219), VoxelUpdateQueue (:221-304).
//
// Tie order inside VoxelUpdateQueue: the reference pops equal-priority chunks in the order that
// `priority-queue 2.0.3` produces. That crate is not vendored in the reference tree and no reference
// test pins the order (PARITY UNPINNED); IncreasePQ below restates the crate's heap mechanics, which
// fix the order without reference to hash values. The order only decides WHICH chunks run in a tick
// when max_update_rate binds; the result of updating a given list is order-independent.
#pragma once
#include <cstdint>
#include <deque>
#include <utility>
#include <unordered_map>
#include <unordered_set>
#include <vector>

namespace vxb {

struct Key3 {
    uint32_t v[3];
    bool operator==(const Key3& o) const { return v[0] == o.v[0] && v[1] == o.v[1] && v[2] == o.v[2]; }
};
struct Key4 {
    uint32_t v[4];
    bool operator==(const Key4& o) const { return v[0] == o.v[0] && v[1] == o.v[1] && v[2] == o.v[2] && v[3] == o.v[3]; }
};
static inline uint64_t mix_u64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}
struct Key3Hash {
    size_t operator()(const Key3& k) const { return (size_t)mix_u64(((uint64_t)k.v[0] << 42) ^ ((uint64_t)k.v[1] << 21) ^ k.v[2]); }
};
struct Key4Hash {
    size_t operator()(const Key4& k) const {
        return (size_t)mix_u64(mix_u64(((uint64_t)k.v[0] << 32) | k.v[1]) ^ (((uint64_t)k.v[2] << 32) | k.v[3]));
    }
};

// src/utils.rs:58-149 with V = bool and keep_value = |a, b| a || b (voxel_sim_manager.rs:522-525)
class QueueMapForced {
   public:
    bool push(const Key3& key, bool value) {
        auto it = map_.find(key);
        if (it != map_.end()) {
            it->second = it->second || value;
            return false;
        }
        queue_.push_back(key);
        map_.emplace(key, value);
        return true;
    }
    bool pop(Key3& key, bool& value) {
        if (queue_.empty()) return false;
        key = queue_.front();
        queue_.pop_front();
        auto it = map_.find(key);
        if (it == map_.end()) return false;
        value = it->second;
        map_.erase(it);
        return true;
    }
    size_t len() const { return queue_.size(); }
    bool is_empty() const { return queue_.empty(); }

   private:
    std::deque<Key3> queue_;
    std::unordered_map<Key3, bool, Key3Hash> map_;
};

// src/utils.rs:151-219
class QueueSet4 {
   public:
    void push(const Key4& item) {
        if (set_.insert(item).second) queue_.push_back(item);
    }
    bool pop(Key4& item) {
        if (queue_.empty()) return false;
        item = queue_.front();
        queue_.pop_front();
        set_.erase(item);
        return true;
    }
    void extend(const std::vector<Key4>& items) {
        for (const Key4& i : items) push(i);
    }
    size_t len() const { return queue_.size(); }
    bool is_empty() const { return queue_.empty(); }

   private:
    std::deque<Key4> queue_;
    std::unordered_set<Key4, Key4Hash> set_;
};

// One PriorityQueue<[u32;3], i32, RandomState> as VoxelUpdateQueue uses it (utils.rs:222, :240, :250, :285):
// push_increase and pop with the pop ORDER of the `priority-queue` crate, version 2.0.3 (client/Cargo.lock:3029-3036).
//
// The crate's source is not under /root/reference, so this restates its published algorithm (PARITY UNPINNED, no
// reference test holds the order). The crate keeps the items in an insertion-ordered IndexMap and a binary max-heap of
// map indices beside it. The hasher (ahash, seed 0) only serves look-ups; the pop order among equal priorities follows
// from the heap mechanics alone:
//   push of a new item   : appended as the last heap position, then bubbled up while parent < new (strict)
//   push of a known item : priority replaced, bubbled up from where it stands, then sifted down from where it lands
//   push_increase        : a push when the item is new or the priority is strictly greater; otherwise nothing
//   pop                  : the root leaves, the LAST heap position takes its place and is sifted down; a child takes the
//                          place only when strictly greater, the left child wins a tie between the children
// The IndexMap's swap_remove relabels map indices on a pop but never moves a heap position, so a heap of items with a
// position look-up gives the same order as the crate's heap of indices (tests/host_oracle.py keeps the crate's full
// three-array store and tests/test_host_queues.py checks the two against each other on random traffic).
class IncreasePQ {
   public:
    void push_increase(const Key3& item, int32_t priority) {
        auto it = pos_.find(item);
        if (it == pos_.end()) {
            heap_.push_back(Node{item, priority});
            bubble_up(heap_.size() - 1);
        } else if (priority > heap_[it->second].priority) {
            heap_[it->second].priority = priority;
            heapify(bubble_up(it->second));
        }
    }
    bool pop(Key3& item, int32_t& priority) {
        if (heap_.empty()) return false;
        item = heap_[0].item;
        priority = heap_[0].priority;
        pos_.erase(item);
        Node last = heap_.back();
        heap_.pop_back();
        if (!heap_.empty()) {
            heap_[0] = last;
            pos_[last.item] = 0;
            heapify(0);
        }
        return true;
    }
    size_t len() const { return heap_.size(); }
    bool is_empty() const { return heap_.empty(); }

   private:
    struct Node {
        Key3 item;
        int32_t priority;
    };
    // moves the node at position i towards the root past every strictly smaller parent; returns where it lands
    size_t bubble_up(size_t i) {
        Node n = heap_[i];
        while (i > 0) {
            size_t parent = (i - 1) / 2;
            if (!(heap_[parent].priority < n.priority)) break;
            heap_[i] = heap_[parent];
            pos_[heap_[i].item] = (uint32_t)i;
            i = parent;
        }
        heap_[i] = n;
        pos_[n.item] = (uint32_t)i;
        return i;
    }
    void heapify(size_t i) {
        const size_t n = heap_.size();
        if (n <= 1) return;
        for (;;) {
            size_t l = 2 * i + 1, r = l + 1, largest = i;
            int32_t lp = heap_[i].priority;
            if (l < n) {
                if (heap_[l].priority > lp) {
                    largest = l;
                    lp = heap_[l].priority;
                }
                if (r < n && heap_[r].priority > lp) largest = r;
            }
            if (largest == i) return;
            std::swap(heap_[i], heap_[largest]);
            pos_[heap_[i].item] = (uint32_t)i;
            pos_[heap_[largest].item] = (uint32_t)largest;
            i = largest;
        }
    }
    std::vector<Node> heap_;
    std::unordered_map<Key3, uint32_t, Key3Hash> pos_;
};

// src/utils.rs:221-304
class VoxelUpdateQueue {
   public:
    void push_all(const Key3& item) { push_with_priority(item, 0); }  // :236-244
    void push_with_priority(const Key3& item, int32_t priority) {     // :246-254
        for (int p = 0; p < 8; p++) sets_[p].push_increase(item, priority);
    }
    void next_queue_set() {  // :256-264, x fastest
        active_[0] = (active_[0] + 1) % 2;
        if (active_[0] == 0) {
            active_[1] = (active_[1] + 1) % 2;
            if (active_[1] == 0) active_[2] = (active_[2] + 1) % 2;
        }
    }
    void swap_queue_set() {  // :275-282; caller guarantees !is_empty()
        do {
            next_queue_set();
        } while (active().is_empty());
    }
    bool pop(Key3& item, int32_t& priority) { return active().pop(item, priority); }
    void queue_set_idx(uint32_t out[3]) const { out[0] = active_[0]; out[1] = active_[1]; out[2] = active_[2]; }
    size_t len() const {
        size_t n = 0;
        for (int p = 0; p < 8; p++) n += sets_[p].len();
        return n;
    }
    bool is_empty() const {
        for (int p = 0; p < 8; p++)
            if (!sets_[p].is_empty()) return false;
        return true;
    }

   private:
    IncreasePQ& active() { return sets_[active_[0] * 4 + active_[1] * 2 + active_[2]]; }
    IncreasePQ sets_[8];  // [x][y][z]
    uint32_t active_[3] = {0, 0, 0};
};

}  // namespace vxb
