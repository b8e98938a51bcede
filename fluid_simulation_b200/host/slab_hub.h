// slab_hub.h -- the rendezvous a multi-GPU run of the C++ host needs, without MPI or Python: fluid_tunnel --gpus N
// forks one process per GPU BEFORE anything touches CUDA; the parent stays behind as a hub on N socket pairs and
// serves all-gathers (a barrier is an all-gather of nothing, a broadcast keeps rank 0's part).  The data path
// never comes through here: halos move between the GPUs by peer stores / NCCL inside libfluid_cuda, frames go
// from each rank's pinned buffer straight into the shared files.  (The reference has no distributed mode; this
// is the C++ counterpart of fluid_simulation_b200/multi_gpu.py.)
#ifndef FLUID_B200_SLAB_HUB_H
#define FLUID_B200_SLAB_HUB_H

#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>
#include <sys/socket.h>
#include <sys/types.h>
#include <sys/wait.h>
#include <unistd.h>

namespace slab {

inline void write_all(int fd, const void* p, size_t n)
{
    const char* c = (const char*)p;
    while (n) { ssize_t k = ::write(fd, c, n); if (k <= 0) throw std::runtime_error("slab hub: write failed"); c += k; n -= (size_t)k; }
}
inline bool read_all(int fd, void* p, size_t n)
{
    char* c = (char*)p;
    while (n) { ssize_t k = ::read(fd, c, n); if (k <= 0) return false; c += k; n -= (size_t)k; }
    return true;
}

// a rank's end of the hub
class Hub {
    int fd_;
public:
    const int rank, nranks;
    Hub(int fd, int r, int n) : fd_(fd), rank(r), nranks(n) {}
    // every rank contributes `bytes` bytes; out receives nranks * bytes, in rank order
    void allgather(const void* mine, size_t bytes, void* out)
    {
        const uint64_t n = bytes;
        write_all(fd_, &n, sizeof n);
        if (bytes) write_all(fd_, mine, bytes);
        if (bytes && !read_all(fd_, out, bytes * (size_t)nranks)) throw std::runtime_error("slab hub: peer lost");
        if (!bytes) { char ack; if (!read_all(fd_, &ack, 1)) throw std::runtime_error("slab hub: peer lost"); }
    }
    void barrier() { allgather(nullptr, 0, nullptr); }
};

// parent side: serve collectives until every rank has closed its socket; returns the worst exit status
inline int serve(std::vector<int>& fds, std::vector<pid_t>& pids)
{
    const int n = (int)fds.size();
    std::vector<char> all, part;
    for (;;) {
        uint64_t bytes = 0;
        int alive = 0;
        for (int r = 0; r < n; ++r) {
            uint64_t b = 0;
            if (fds[r] >= 0 && read_all(fds[r], &b, sizeof b)) { ++alive; bytes = b; }
            else if (fds[r] >= 0) { ::close(fds[r]); fds[r] = -1; }
        }
        if (alive == 0) break;
        if (alive != n) break;                            // a rank died in the middle: let the others fail on read
        all.resize(bytes * (size_t)n);
        for (int r = 0; r < n; ++r) if (bytes && !read_all(fds[r], all.data() + bytes * (size_t)r, bytes)) { ::close(fds[r]); fds[r] = -1; }
        for (int r = 0; r < n; ++r) {
            if (fds[r] < 0) continue;
            if (bytes) write_all(fds[r], all.data(), all.size()); else { char ack = 1; write_all(fds[r], &ack, 1); }
        }
    }
    for (int r = 0; r < n; ++r) if (fds[r] >= 0) ::close(fds[r]);
    int worst = 0;
    for (pid_t p : pids) { int st = 0; ::waitpid(p, &st, 0); const int rc = WIFEXITED(st) ? WEXITSTATUS(st) : 128; if (rc > worst) worst = rc; }
    return worst;
}

// fork `n` ranks; in a child returns its Hub (heap-allocated), in the parent serves and exits the process
inline Hub* spawn(int n)
{
    std::vector<int> fds((size_t)n, -1); std::vector<pid_t> pids;
    for (int r = 0; r < n; ++r) {
        int sv[2];
        if (::socketpair(AF_UNIX, SOCK_STREAM, 0, sv) != 0) throw std::runtime_error("socketpair failed");
        const pid_t p = ::fork();
        if (p < 0) throw std::runtime_error("fork failed");
        if (p == 0) {                                     // child: keep only its own end
            ::close(sv[0]);
            for (int q = 0; q < r; ++q) ::close(fds[q]);
            return new Hub(sv[1], r, n);
        }
        ::close(sv[1]); fds[(size_t)r] = sv[0]; pids.push_back(p);
    }
    ::_exit(serve(fds, pids));
}

}  // namespace slab
#endif
