// shm_allreduce.cpp — cross-process gradient all-reduce over POSIX shared memory, fused with
// the 1/num_workers averaging and the SGD update (CPU backend of the strategy).
//
// This is the CPU stand-in for the NVLink peer-memory kernels: every worker process on the
// host maps one segment holding a slot per rank plus a result buffer.
//   1. each rank publishes its gradient arena into its slot           (barrier)
//   2. rank r sums shard r over all slots in rank order -> result     (barrier)
//   3. every rank applies w -= lr/N * result (tail slots: w += result) from the same bytes,
//      so replicas stay bit-identical — the "mirrored" contract (README.md:364, 229-231).
// The reference's equivalent is TensorFlow's collective ring all-reduce over gRPC
// (README.md:395-403); none of it is reproduced here.
#include <errno.h>
#include <fcntl.h>
#include <sched.h>
#include <signal.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>

#define B2C_API extern "C" __attribute__((visibility("default")))

namespace {

struct alignas(64) Header {
  std::atomic<uint32_t> magic;
  std::atomic<uint32_t> attached;
  std::atomic<uint32_t> arrive;   // barrier arrivals in the current generation
  std::atomic<uint32_t> generation;
  int32_t world;
  int64_t capacity;  // floats per slot
  int32_t creator_pid;   // rank 0 of THIS run: attachers reject a segment whose creator is gone (left by a crashed run)
  int64_t created_ns;    // CLOCK_REALTIME at creation, for the error message
};

constexpr uint32_t kMagic = 0xB200D157u;

thread_local char t_err[256] = "";

}  // namespace

struct b2c_shm_comm {
  Header* hdr = nullptr;
  float* slots = nullptr;   // [world][capacity]
  float* result = nullptr;  // [capacity]
  size_t bytes = 0;
  int rank = 0, world = 1;
  int64_t capacity = 0;
  std::string name;
  double timeout_s = 120.0;
};

static double now_s() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec + ts.tv_nsec * 1e-9;
}

B2C_API const char* b2c_last_error(void) { return t_err; }

// Sense-free generation barrier; fails (returns -1) instead of hanging forever when a peer
// died — the fail-stop behaviour the reference describes (README.md:400).
static int barrier(b2c_shm_comm* c) {
  Header* h = c->hdr;
  const uint32_t gen = h->generation.load(std::memory_order_acquire);
  if (h->arrive.fetch_add(1, std::memory_order_acq_rel) + 1 == (uint32_t)c->world) {
    h->arrive.store(0, std::memory_order_relaxed);
    h->generation.store(gen + 1, std::memory_order_release);
    return 0;
  }
  const double t0 = now_s();
  int spins = 0;
  while (h->generation.load(std::memory_order_acquire) == gen) {
    if (++spins > 200) {
      sched_yield();
      if ((spins & 0x3ff) == 0 && now_s() - t0 > c->timeout_s) {
        snprintf(t_err, sizeof(t_err), "shm barrier timed out after %.0f s (rank %d of %d): a peer is gone",
                 c->timeout_s, c->rank, c->world);
        return -1;
      }
    }
  }
  return 0;
}

B2C_API b2c_shm_comm* b2c_shm_create(const char* name, int rank, int world, int64_t capacity_floats,
                                     double timeout_s) {
  if (world < 1 || rank < 0 || rank >= world || capacity_floats <= 0) {
    snprintf(t_err, sizeof(t_err), "shm_create: bad arguments");
    return nullptr;
  }
  b2c_shm_comm* c = new b2c_shm_comm();
  c->rank = rank, c->world = world, c->capacity = capacity_floats, c->name = name;
  c->timeout_s = timeout_s > 0 ? timeout_s : 120.0;
  c->bytes = sizeof(Header) + sizeof(float) * (size_t)capacity_floats * (world + 1);
  int fd = -1;
  void* p = MAP_FAILED;
  const double t0 = now_s();
  if (rank == 0) {
    shm_unlink(name);
    fd = shm_open(name, O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd >= 0 && ftruncate(fd, (off_t)c->bytes) != 0) {
      close(fd);
      fd = -1;
    }
    if (fd < 0) {
      snprintf(t_err, sizeof(t_err), "shm_create: cannot create %s: %s", name, strerror(errno));
      delete c;
      return nullptr;
    }
    p = mmap(nullptr, c->bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (p == MAP_FAILED) {
      snprintf(t_err, sizeof(t_err), "shm_create: mmap failed: %s", strerror(errno));
      delete c;
      return nullptr;
    }
    c->hdr = static_cast<Header*>(p);
    timespec ts;
    clock_gettime(CLOCK_REALTIME, &ts);
    c->hdr->world = world;
    c->hdr->capacity = capacity_floats;
    c->hdr->creator_pid = static_cast<int32_t>(getpid());
    c->hdr->created_ns = static_cast<int64_t>(ts.tv_sec) * 1000000000ll + ts.tv_nsec;
    c->hdr->arrive.store(0);
    c->hdr->generation.store(0);
    c->hdr->attached.store(1);
    c->hdr->magic.store(kMagic, std::memory_order_release);
  } else {
    // Open, map, wait for rank 0's magic — and make sure the segment belongs to a LIVE run: the name repeats
    // across runs (TF_CONFIG ports are fixed), so a segment left behind by a crashed run may still be there until
    // this run's rank 0 replaces it.  A segment whose creator process no longer exists is dropped and reopened.
    while (true) {
      fd = shm_open(name, O_RDWR, 0600);
      if (fd >= 0) {
        struct stat st;
        if (fstat(fd, &st) == 0 && (size_t)st.st_size >= c->bytes) {
          p = mmap(nullptr, c->bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
          close(fd);
          fd = -1;
          if (p != MAP_FAILED) {
            Header* h = static_cast<Header*>(p);
            bool stale = false;
            while (h->magic.load(std::memory_order_acquire) != kMagic && now_s() - t0 <= c->timeout_s) usleep(1000);
            if (h->magic.load(std::memory_order_acquire) == kMagic) {
              const pid_t creator = static_cast<pid_t>(h->creator_pid);
              stale = creator <= 0 || (kill(creator, 0) != 0 && errno == ESRCH) || h->world != world ||
                      h->capacity != capacity_floats;
              if (!stale) {
                c->hdr = h;
                break;
              }
            }
            munmap(p, c->bytes);
            p = MAP_FAILED;
          }
        } else {
          close(fd);
          fd = -1;
        }
      }
      if (now_s() - t0 > c->timeout_s) break;
      usleep(2000);
    }
    if (c->hdr == nullptr) {
      snprintf(t_err, sizeof(t_err), "shm_create: no live segment %s appeared within %.0f s (rank 0 missing, or only "
               "a stale segment of a crashed run was found)", name, c->timeout_s);
      delete c;
      return nullptr;
    }
    c->hdr->attached.fetch_add(1);
  }
  c->slots = reinterpret_cast<float*>(reinterpret_cast<char*>(c->hdr) + sizeof(Header));
  c->result = c->slots + (size_t)capacity_floats * world;
  if (barrier(c) != 0) {  // everyone is attached: rank 0 may unlink the name
    munmap(c->hdr, c->bytes);
    if (rank == 0) shm_unlink(name);
    delete c;
    return nullptr;
  }
  if (rank == 0) shm_unlink(name);
  return c;
}

B2C_API void b2c_shm_destroy(b2c_shm_comm* c) {
  if (c == nullptr) return;
  if (c->hdr != nullptr) munmap(c->hdr, c->bytes);
  delete c;
}

B2C_API int b2c_shm_barrier(b2c_shm_comm* c) { return barrier(c); }

static int reduce_into_result(b2c_shm_comm* c, const float* src, int64_t n) {
  if (n > c->capacity) {
    snprintf(t_err, sizeof(t_err), "shm all-reduce: %lld floats exceed capacity %lld", (long long)n,
             (long long)c->capacity);
    return -1;
  }
  std::memcpy(c->slots + (size_t)c->rank * c->capacity, src, sizeof(float) * n);
  if (barrier(c) != 0) return -1;
  const int64_t per = (n + c->world - 1) / c->world;
  const int64_t b = std::min<int64_t>(n, per * c->rank), e = std::min<int64_t>(n, b + per);
  for (int64_t i = b; i < e; ++i) {
    float s = 0.f;
    for (int r = 0; r < c->world; ++r) s += c->slots[(size_t)r * c->capacity + i];  // fixed rank order
    c->result[i] = s;
  }
  return barrier(c);
}

// Plain sum all-reduce, in place.
B2C_API int b2c_shm_allreduce(b2c_shm_comm* c, float* buf, int64_t n) {
  if (reduce_into_result(c, buf, n) != 0) return -1;
  std::memcpy(buf, c->result, sizeof(float) * n);
  return barrier(c);  // result may be overwritten by the next call only after everyone copied
}

// Fused gradient all-reduce + 1/world averaging + SGD; tail slots [n_params, n_total) are
// summed into w unscaled (loss / accuracy counters); g is zeroed.            [K11+K12+K13]
B2C_API int b2c_shm_allreduce_sgd(b2c_shm_comm* c, float* w, float* g, int64_t n_params,
                                  int64_t n_total, float lr) {
  if (reduce_into_result(c, g, n_total) != 0) return -1;
  const float scale = lr / (float)c->world;
  for (int64_t i = 0; i < n_params; ++i) w[i] -= scale * c->result[i];
  for (int64_t i = n_params; i < n_total; ++i) w[i] += c->result[i];
  std::memset(g, 0, sizeof(float) * n_total);
  return barrier(c);
}
