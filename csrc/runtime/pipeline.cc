// Host-side input pipeline runtime (the role tf.data's C++ runtime plays for the reference:
// `from_tensor_slices(...).shuffle(B).repeat().batch(N).prefetch(P)`, mnist_keras_distributed.py:142-145,
// distributed_with_keras.py:30,54).
//
// One background thread per pipeline produces batches into a ring of `depth` slots:
//   index stream:  [shard]  ->  windowed shuffle (tf.data semantics: fill a buffer of `shuffle_buffer`
//                  indices, emit a uniformly chosen one, refill; drained and reshuffled at every epoch)
//                  ->  repeat(count | forever)
//   batch:         gather `batch` rows of every component (features, labels) into the slot
// The consumer (`tfde_pipe_next_into`) copies a ready slot into caller memory -- typically a pinned host
// tensor from PyTorch's caching host allocator, which is what makes the subsequent non-blocking H2D copy
// safe to overlap -- and hands the slot back. No CUDA dependency: builds and runs on CPU-only boxes.
#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

#define TFDE_API extern "C" __attribute__((visibility("default")))

namespace {

struct SplitMix64 {
  uint64_t s;
  explicit SplitMix64(uint64_t seed) : s(seed) {}
  uint64_t next() {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  }
  // unbiased enough for shuffling (Lemire reduction)
  uint64_t below(uint64_t n) { return (uint64_t)(((__uint128_t)next() * n) >> 64); }
};

struct Slot {
  std::vector<uint8_t> feat, label;
  int64_t count = 0;
  bool ready = false;
};

class Runner {
 public:
  Runner(const void* feat, int64_t feat_bytes, const void* label, int64_t label_bytes, int64_t n,
         int64_t shuffle_buffer, uint64_t seed, int64_t repeat, int64_t batch, int drop_remainder, int depth,
         int shard_num, int shard_index)
      : feat_((const uint8_t*)feat), label_((const uint8_t*)label), fb_(feat_bytes), lb_(label ? label_bytes : 0),
        n_(n), shuf_(shuffle_buffer), seed_(seed), repeat_(repeat), batch_(batch), drop_(drop_remainder != 0),
        shard_num_(shard_num < 1 ? 1 : shard_num), shard_index_(shard_index), ring_((size_t)(depth < 2 ? 2 : depth)) {
    for (auto& s : ring_) {
      s.feat.resize((size_t)(batch_ * fb_));
      s.label.resize((size_t)(batch_ * lb_));
    }
    worker_ = std::thread([this] { produce(); });
  }
  ~Runner() {
    {
      std::lock_guard<std::mutex> lk(mu_);
      stop_ = true;
    }
    cv_free_.notify_all();
    cv_ready_.notify_all();
    if (worker_.joinable()) worker_.join();
  }

  // Blocks until a batch is ready; copies it out; returns its row count (0 == end of data).
  int64_t next_into(void* feat_dst, void* label_dst) {
    std::unique_lock<std::mutex> lk(mu_);
    cv_ready_.wait(lk, [this] { return ring_[tail_].ready || finished_ || stop_; });
    if (!ring_[tail_].ready) return 0;
    Slot& s = ring_[tail_];
    const int64_t c = s.count;
    lk.unlock();   // the slot is owned by the consumer until `ready` is cleared
    std::memcpy(feat_dst, s.feat.data(), (size_t)(c * fb_));
    if (lb_ && label_dst) std::memcpy(label_dst, s.label.data(), (size_t)(c * lb_));
    lk.lock();
    s.ready = false;
    tail_ = (tail_ + 1) % ring_.size();
    lk.unlock();
    cv_free_.notify_one();
    return c;
  }

 private:
  bool push_batch(const std::vector<int64_t>& idx) {
    std::unique_lock<std::mutex> lk(mu_);
    cv_free_.wait(lk, [this] { return !ring_[head_].ready || stop_; });
    if (stop_) return false;
    Slot& s = ring_[head_];
    lk.unlock();
    for (size_t r = 0; r < idx.size(); ++r) {
      std::memcpy(s.feat.data() + r * (size_t)fb_, feat_ + (size_t)idx[r] * (size_t)fb_, (size_t)fb_);
      if (lb_) std::memcpy(s.label.data() + r * (size_t)lb_, label_ + (size_t)idx[r] * (size_t)lb_, (size_t)lb_);
    }
    lk.lock();
    s.count = (int64_t)idx.size();
    s.ready = true;
    head_ = (head_ + 1) % ring_.size();
    lk.unlock();
    cv_ready_.notify_one();
    return true;
  }

  void produce() {
    std::vector<int64_t> base;
    for (int64_t i = shard_index_; i < n_; i += shard_num_) base.push_back(i);
    std::vector<int64_t> buf, cur;
    cur.reserve((size_t)batch_);
    bool alive = true;
    auto emit = [&](int64_t v) {
      cur.push_back(v);
      if ((int64_t)cur.size() == batch_) {
        alive = push_batch(cur);
        cur.clear();
      }
    };
    int64_t epoch = 0;
    while (alive && !base.empty() && (repeat_ < 0 || epoch < repeat_)) {
      if (shuf_ > 1) {
        SplitMix64 rng(seed_ + 0x632BE59BD9B4E019ull * (uint64_t)(epoch + 1));
        buf.clear();
        for (size_t i = 0; i < base.size() && alive; ++i) {
          if ((int64_t)buf.size() < shuf_) { buf.push_back(base[i]); continue; }
          const size_t j = (size_t)rng.below(buf.size());
          const int64_t out = buf[j];
          buf[j] = base[i];
          emit(out);
        }
        while (!buf.empty() && alive) {
          const size_t j = (size_t)rng.below(buf.size());
          const int64_t out = buf[j];
          buf[j] = buf.back();
          buf.pop_back();
          emit(out);
        }
      } else {
        for (size_t i = 0; i < base.size() && alive; ++i) emit(base[i]);
      }
      ++epoch;
    }
    if (alive && !cur.empty() && !drop_) push_batch(cur);
    {
      std::lock_guard<std::mutex> lk(mu_);
      finished_ = true;
    }
    cv_ready_.notify_all();
  }

  const uint8_t* feat_; const uint8_t* label_;
  int64_t fb_, lb_, n_, shuf_; uint64_t seed_; int64_t repeat_, batch_; bool drop_;
  int shard_num_, shard_index_;
  std::vector<Slot> ring_;
  size_t head_ = 0, tail_ = 0;
  bool finished_ = false, stop_ = false;
  std::mutex mu_;
  std::condition_variable cv_ready_, cv_free_;
  std::thread worker_;
};

}  // namespace

// features/labels: contiguous arrays of n rows; the caller keeps them alive for the pipeline's lifetime.
// repeat < 0: forever. shuffle_buffer <= 1: no shuffling. depth: prefetch ring size (>= 2).
TFDE_API void* tfde_pipe_create(const void* features, int64_t feat_bytes, const void* labels, int64_t label_bytes,
                                int64_t n, int64_t shuffle_buffer, uint64_t seed, int64_t repeat, int64_t batch,
                                int drop_remainder, int depth, int shard_num, int shard_index) {
  if (!features || feat_bytes <= 0 || n < 0 || batch <= 0) return nullptr;
  return new Runner(features, feat_bytes, labels, label_bytes, n, shuffle_buffer, seed, repeat, batch, drop_remainder,
                    depth, shard_num, shard_index);
}
TFDE_API int64_t tfde_pipe_next_into(void* pipe, void* feat_dst, void* label_dst) {
  return static_cast<Runner*>(pipe)->next_into(feat_dst, label_dst);
}
TFDE_API void tfde_pipe_destroy(void* pipe) { delete static_cast<Runner*>(pipe); }
TFDE_API int tfde_runtime_abi_version() { return 1; }
