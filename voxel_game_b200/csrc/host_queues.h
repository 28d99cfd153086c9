// Host work queues of the compute pass, re-expressed in C++.
// Mirrors src/utils.rs: QueueMap (:58-149), QueueSet (:151-219), VoxelUpdateQueue (:221-304).
//
// Tie order inside VoxelUpdateQueue: the reference pops equal-priority chunks in the iteration
// order of `priority_queue 2.0.3` over `ahash 0.8.11` (seed 0); neither crate is vendored in the
// reference tree and no reference test pins the order, so it is UNPINNED. This implementation is
// deterministic instead: among equal priorities the entry whose priority was set first pops first.
// The order only decides WHICH chunks run in a tick when max_update_rate binds; the result of
// updating a given list is order-independent.
#pragma once
#include <cstdint>
#include <deque>
#include <queue>
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

// One PriorityQueue<[u32;3], i32> with push_increase / pop (priority-queue crate semantics):
// push_increase inserts, or raises the priority of an existing item, never lowers it.
class IncreasePQ {
   public:
    void push_increase(const Key3& item, int32_t priority) {
        auto it = prio_.find(item);
        if (it == prio_.end()) {
            prio_.emplace(item, priority);
            heap_.push(Entry{priority, seq_++, item});
        } else if (priority > it->second) {
            it->second = priority;
            heap_.push(Entry{priority, seq_++, item});  // the stale lower entry is skipped on pop
        }
    }
    bool pop(Key3& item, int32_t& priority) {
        while (!heap_.empty()) {
            Entry e = heap_.top();
            heap_.pop();
            auto it = prio_.find(e.item);
            if (it != prio_.end() && it->second == e.priority) {
                item = e.item;
                priority = e.priority;
                prio_.erase(it);
                return true;
            }
        }
        return false;
    }
    size_t len() const { return prio_.size(); }
    bool is_empty() const { return prio_.empty(); }

   private:
    struct Entry {
        int32_t priority;
        uint64_t seq;
        Key3 item;
        bool operator<(const Entry& o) const {  // max-heap on priority, then FIFO
            if (priority != o.priority) return priority < o.priority;
            return seq > o.seq;
        }
    };
    std::unordered_map<Key3, int32_t, Key3Hash> prio_;
    std::priority_queue<Entry> heap_;
    uint64_t seq_ = 0;
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
