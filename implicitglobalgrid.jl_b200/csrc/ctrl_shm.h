// Host-side control plane between the ranks of one box: one POSIX shared-memory mailbox per rank.
//
// The halo data path never touches it.  It is used for the rare collective host steps the library must do on its
// own after the caller-driven bootstrap: exchanging new CUDA-IPC handles when a staging or gather buffer grows,
// and the barrier of gather!.  It stands in for the MPI calls the reference would make at those points
// (MPI.Barrier, ref: src/tools.jl:896-897; the implicit synchronisation of MPI.Gatherv!, ref: src/gather.jl:49).
#pragma once

#include <atomic>
#include <cstdint>
#include <string>
#include <vector>

namespace igg {

constexpr int kMailPayload = 320;

struct Mailbox {
    std::atomic<uint64_t> pub_seq;     // sequence number of the last published payload
    std::atomic<uint64_t> bar_seq;     // sequence number of the last barrier this rank entered
    char payload[2][kMailPayload];     // double-buffered by sequence parity
};

class ControlPlane {
public:
    ~ControlPlane();
    // Creates my mailbox; `name_out` is what travels in the bootstrap blob.
    int create(int me, int nprocs, std::string& name_out);
    // Maps the mailboxes of all ranks (names in rank order).
    int attach(const std::vector<std::string>& names);
    // Every rank contributes kMailPayload bytes, everyone gets all of them (rank order).  Collective.
    int allgather(const void* mine, std::vector<char>& all);
    int barrier();
    void close();
    bool attached() const { return !boxes_.empty(); }

private:
    int wait_ge(std::atomic<uint64_t>& a, uint64_t v);
    int me_ = 0, nprocs_ = 1;
    std::string my_name_;
    Mailbox* mine_ = nullptr;
    std::vector<Mailbox*> boxes_;
    uint64_t seq_ = 0, bseq_ = 0;
};

}  // namespace igg
