// rrt_hostframe.cu — one canvas for the one-process-per-GPU ranks of a node, in HOST memory.
//
// The reference's output is `Canvas.buffer: Vec<u32>` in host memory (engine.rs:123-128, 146-158).  With the frame
// sharded over N GPUs (interleaved tile rows) the shortest way there is N device->host copies in
// parallel, one PCIe link each, straight into ONE page-locked frame every rank maps: a POSIX shared-memory segment
// = one header page + width*height words.  The header also carries what a host without torch.distributed needs to
// run the ranks: a process-shared barrier and a 64-byte slot for the CUDA IPC handle of the peer-store mode
// (rrt_render_device_into_frame).  No counterpart in the reference (rayon threads share one address space).
#include <atomic>
#include <cerrno>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>

#include <cuda_runtime.h>
#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include "../../include/rrt.h"

extern "C" void rrt_set_last_error_(const char* msg);

namespace {

constexpr uint64_t kMagic = 0x52525446524D3031ull;   // "RRTFRM01"
constexpr size_t kHeaderBytes = 4096;

struct Header {
    std::atomic<uint64_t> magic;          // written LAST by the creator: an opener waits for it
    int32_t width, height, n_ranks, _pad;
    std::atomic<uint32_t> arrived, generation;
    std::atomic<uint32_t> ipc_seq;        // bumped by rrt_host_frame_publish_handle
    unsigned char ipc[64];
};
static_assert(sizeof(Header) <= kHeaderBytes, "header page");
static_assert(std::atomic<uint32_t>::is_always_lock_free && std::atomic<uint64_t>::is_always_lock_free, "shm atomics");

int fail(int code, const std::string& msg) {
    rrt_set_last_error_(msg.c_str());
    return code;
}

double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

}  // namespace

struct rrt_host_frame {
    std::string name;
    void* base = nullptr;
    size_t bytes = 0;
    bool creator = false, pinned = false;
    int pin_device = -1;
    uint32_t ipc_seen = 0;
    Header* hdr() const { return static_cast<Header*>(base); }
    uint32_t* pixels() const { return reinterpret_cast<uint32_t*>(static_cast<char*>(base) + kHeaderBytes); }
};

extern "C" {

int rrt_host_frame_create(const char* name, int32_t width, int32_t height, int32_t n_ranks, rrt_host_frame** out) {
    if (!name || name[0] != '/' || !out || width <= 0 || height <= 0 || n_ranks < 1)
        return fail(RRT_ERR_INVALID, "rrt_host_frame_create: name must start with '/', sizes positive");
    if ((int64_t)width * height > (int64_t)1 << 30) return fail(RRT_ERR_INVALID, "frame too large");
    *out = nullptr;
    shm_unlink(name);   // a stale segment of a crashed run
    const int fd = shm_open(name, O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0) return fail(RRT_ERR_IO, std::string("shm_open(create): ") + std::strerror(errno));
    const size_t bytes = kHeaderBytes + (size_t)width * height * sizeof(uint32_t);
    if (ftruncate(fd, (off_t)bytes) != 0) {
        const std::string e = std::strerror(errno);
        close(fd);
        shm_unlink(name);
        return fail(RRT_ERR_IO, "ftruncate: " + e);
    }
    void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (p == MAP_FAILED) {
        shm_unlink(name);
        return fail(RRT_ERR_IO, std::string("mmap: ") + std::strerror(errno));
    }
    auto* f = new rrt_host_frame;
    f->name = name; f->base = p; f->bytes = bytes; f->creator = true;
    Header* h = f->hdr();   // fresh segment: zero-filled by the kernel (row 0 of the canvas stays 0, engine.rs:146-158)
    h->width = width; h->height = height; h->n_ranks = n_ranks;
    h->arrived.store(0); h->generation.store(0); h->ipc_seq.store(0);
    h->magic.store(kMagic, std::memory_order_release);
    *out = f;
    return RRT_OK;
}

int rrt_host_frame_open(const char* name, double timeout_s, rrt_host_frame** out) {
    if (!name || !out) return fail(RRT_ERR_INVALID, "null argument");
    *out = nullptr;
    const double t_end = now_s() + (timeout_s > 0 ? timeout_s : 60.0);
    int fd = -1;
    struct stat st {};
    for (;;) {   // the creator may not be there yet
        fd = shm_open(name, O_RDWR, 0600);
        if (fd >= 0 && fstat(fd, &st) == 0 && (size_t)st.st_size > kHeaderBytes) break;
        if (fd >= 0) close(fd);
        fd = -1;
        if (now_s() > t_end) return fail(RRT_ERR_IO, std::string("rrt_host_frame_open: no segment ") + name);
        std::this_thread::sleep_for(std::chrono::milliseconds(2));
    }
    void* p = mmap(nullptr, (size_t)st.st_size, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (p == MAP_FAILED) return fail(RRT_ERR_IO, std::string("mmap: ") + std::strerror(errno));
    auto* f = new rrt_host_frame;
    f->name = name; f->base = p; f->bytes = (size_t)st.st_size;
    while (f->hdr()->magic.load(std::memory_order_acquire) != kMagic) {
        if (now_s() > t_end) {
            munmap(p, f->bytes);
            delete f;
            return fail(RRT_ERR_IO, "rrt_host_frame_open: segment never initialised");
        }
        std::this_thread::sleep_for(std::chrono::milliseconds(1));
    }
    const Header* h = f->hdr();
    if (kHeaderBytes + (size_t)h->width * h->height * sizeof(uint32_t) != f->bytes) {
        munmap(p, f->bytes);
        delete f;
        return fail(RRT_ERR_IO, "rrt_host_frame_open: size mismatch");
    }
    *out = f;
    return RRT_OK;
}

uint32_t* rrt_host_frame_pixels(rrt_host_frame* f) { return f ? f->pixels() : nullptr; }

int rrt_host_frame_geometry(const rrt_host_frame* f, int32_t* width, int32_t* height, int32_t* n_ranks) {
    if (!f) return fail(RRT_ERR_INVALID, "null frame");
    if (width) *width = f->hdr()->width;
    if (height) *height = f->hdr()->height;
    if (n_ranks) *n_ranks = f->hdr()->n_ranks;
    return RRT_OK;
}

int rrt_host_frame_pin(rrt_host_frame* f, int device) {
    if (!f) return fail(RRT_ERR_INVALID, "null frame");
    if (f->pinned) return RRT_OK;
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaHostRegister(f->pixels(), f->bytes - kHeaderBytes, cudaHostRegisterPortable);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(RRT_ERR_CUDA, std::string("cudaHostRegister: ") + cudaGetErrorString(e));
    }
    f->pinned = true;
    f->pin_device = device;
    return RRT_OK;
}

// Sense-reversing barrier over the n_ranks processes that mapped the segment.  Never waits forever: a rank that
// died would otherwise hang its peers (and the GPU box).
int rrt_host_frame_barrier(rrt_host_frame* f, double timeout_s) {
    if (!f) return fail(RRT_ERR_INVALID, "null frame");
    Header* h = f->hdr();
    const uint32_t n = (uint32_t)h->n_ranks;
    if (n <= 1) return RRT_OK;
    const uint32_t gen = h->generation.load(std::memory_order_acquire);
    if (h->arrived.fetch_add(1, std::memory_order_acq_rel) + 1 == n) {
        h->arrived.store(0, std::memory_order_relaxed);
        h->generation.store(gen + 1, std::memory_order_release);
        return RRT_OK;
    }
    const double t_end = now_s() + (timeout_s > 0 ? timeout_s : 120.0);
    for (uint32_t spins = 0; h->generation.load(std::memory_order_acquire) == gen; ++spins) {
        if ((spins & 63u) == 63u) {
            sched_yield();
            if ((spins & 0xFFFFu) == 0xFFFFu && now_s() > t_end) return fail(RRT_ERR_IO, "rrt_host_frame_barrier: timeout");
        }
    }
    return RRT_OK;
}

int rrt_host_frame_publish_handle(rrt_host_frame* f, const unsigned char handle[64]) {
    if (!f || !handle) return fail(RRT_ERR_INVALID, "null argument");
    std::memcpy(f->hdr()->ipc, handle, 64);
    f->hdr()->ipc_seq.fetch_add(1, std::memory_order_release);
    return RRT_OK;
}

int rrt_host_frame_fetch_handle(rrt_host_frame* f, double timeout_s, unsigned char handle[64]) {
    if (!f || !handle) return fail(RRT_ERR_INVALID, "null argument");
    const double t_end = now_s() + (timeout_s > 0 ? timeout_s : 60.0);
    while (f->hdr()->ipc_seq.load(std::memory_order_acquire) == f->ipc_seen) {
        if (now_s() > t_end) return fail(RRT_ERR_IO, "rrt_host_frame_fetch_handle: timeout");
        std::this_thread::sleep_for(std::chrono::milliseconds(1));
    }
    f->ipc_seen = f->hdr()->ipc_seq.load(std::memory_order_acquire);
    std::memcpy(handle, f->hdr()->ipc, 64);
    return RRT_OK;
}

void rrt_host_frame_close(rrt_host_frame* f) {
    if (!f) return;
    if (f->pinned) {
        cudaSetDevice(f->pin_device);
        cudaHostUnregister(f->pixels());
        cudaGetLastError();
    }
    munmap(f->base, f->bytes);
    if (f->creator) shm_unlink(f->name.c_str());
    delete f;
}

}  // extern "C"
