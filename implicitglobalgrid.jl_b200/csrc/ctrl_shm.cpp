#include "ctrl_shm.h"

#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include <cstdlib>
#include <cstring>

#include "igg_internal.h"

namespace igg {

static double now_s() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

static double host_timeout_s() {
    const char* e = getenv("IGG_HOST_TIMEOUT_S");
    return e ? atof(e) : 120.0;
}

ControlPlane::~ControlPlane() { close(); }

int ControlPlane::create(int me, int nprocs, std::string& name_out) {
    me_ = me;
    nprocs_ = nprocs;
    uint64_t nonce = 0;
    int fd = ::open("/dev/urandom", O_RDONLY);
    if (fd >= 0) { if (::read(fd, &nonce, sizeof(nonce)) != (ssize_t)sizeof(nonce)) nonce = 0; ::close(fd); }
    if (!nonce) nonce = (uint64_t)getpid() * 2654435761u ^ (uint64_t)(now_s() * 1e9);
    char name[64];
    snprintf(name, sizeof(name), "/igg_b200_%016llx_%d", (unsigned long long)nonce, me);
    fd = shm_open(name, O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0) return fail(IGG_ERR_IPC, "shm_open(%s) failed: %s", name, strerror(errno));
    if (ftruncate(fd, sizeof(Mailbox)) != 0) { ::close(fd); shm_unlink(name); return fail(IGG_ERR_IPC, "ftruncate(%s) failed: %s", name, strerror(errno)); }
    void* p = mmap(nullptr, sizeof(Mailbox), PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    ::close(fd);
    if (p == MAP_FAILED) { shm_unlink(name); return fail(IGG_ERR_IPC, "mmap(%s) failed: %s", name, strerror(errno)); }
    std::memset(p, 0, sizeof(Mailbox));
    mine_ = static_cast<Mailbox*>(p);
    my_name_ = name;
    name_out = name;
    return IGG_OK;
}

int ControlPlane::attach(const std::vector<std::string>& names) {
    if ((int)names.size() != nprocs_) return fail(IGG_ERR_ARG, "control plane: expected %d mailbox names.", nprocs_);
    boxes_.assign(nprocs_, nullptr);
    for (int r = 0; r < nprocs_; ++r) {
        if (r == me_) { boxes_[r] = mine_; continue; }
        int fd = shm_open(names[r].c_str(), O_RDWR, 0600);
        if (fd < 0) return fail(IGG_ERR_IPC, "shm_open(%s) of rank %d failed: %s", names[r].c_str(), r, strerror(errno));
        void* p = mmap(nullptr, sizeof(Mailbox), PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
        ::close(fd);
        if (p == MAP_FAILED) return fail(IGG_ERR_IPC, "mmap(%s) failed: %s", names[r].c_str(), strerror(errno));
        boxes_[r] = static_cast<Mailbox*>(p);
    }
    // Everyone has mapped everyone: the names can go away (the mappings stay valid).
    int rc = barrier();
    if (rc) return rc;
    if (!my_name_.empty()) { shm_unlink(my_name_.c_str()); my_name_.clear(); }
    return IGG_OK;
}

int ControlPlane::wait_ge(std::atomic<uint64_t>& a, uint64_t v) {
    const double t0 = now_s(), limit = host_timeout_s();
    unsigned spins = 0;
    while (a.load(std::memory_order_acquire) < v) {
        if ((++spins & 0xfff) == 0) {
            sched_yield();
            if (now_s() - t0 > limit) return fail(IGG_ERR_TIMEOUT, "control plane: a rank did not arrive within %.0f s.", limit);
        }
    }
    return IGG_OK;
}

int ControlPlane::barrier() {
    if (nprocs_ == 1) return IGG_OK;
    if (boxes_.empty()) return fail(IGG_ERR_STATE, "control plane used before igg_bootstrap_import.");
    ++bseq_;
    mine_->bar_seq.store(bseq_, std::memory_order_release);
    for (int r = 0; r < nprocs_; ++r) {
        int rc = wait_ge(boxes_[r]->bar_seq, bseq_);
        if (rc) return rc;
    }
    return IGG_OK;
}

int ControlPlane::allgather(const void* mine, std::vector<char>& all) {
    all.resize((size_t)nprocs_ * kMailPayload);
    if (nprocs_ == 1) { std::memcpy(all.data(), mine, kMailPayload); return IGG_OK; }
    if (boxes_.empty()) return fail(IGG_ERR_STATE, "control plane used before igg_bootstrap_import.");
    ++seq_;
    std::memcpy(mine_->payload[seq_ & 1], mine, kMailPayload);
    mine_->pub_seq.store(seq_, std::memory_order_release);
    for (int r = 0; r < nprocs_; ++r) {
        int rc = wait_ge(boxes_[r]->pub_seq, seq_);
        if (rc) return rc;
        std::memcpy(all.data() + (size_t)r * kMailPayload, boxes_[r]->payload[seq_ & 1], kMailPayload);
    }
    return barrier();
}

void ControlPlane::close() {
    for (int r = 0; r < (int)boxes_.size(); ++r)
        if (boxes_[r] && boxes_[r] != mine_) munmap(boxes_[r], sizeof(Mailbox));
    boxes_.clear();
    if (mine_) { munmap(mine_, sizeof(Mailbox)); mine_ = nullptr; }
    if (!my_name_.empty()) { shm_unlink(my_name_.c_str()); my_name_.clear(); }
}

}  // namespace igg

// ---- C ABI: control-plane self test (host only; lets the CPU test-suite exercise the mailbox across real
// processes without a GPU) ------------------------------------------------------------------------------------
extern "C" {

int igg_ctrl_selftest_create(int me, int nprocs, char name_out[64], void** handle) {
    if (!name_out || !handle || nprocs < 1 || me < 0 || me >= nprocs) return igg::fail(IGG_ERR_ARG, "igg_ctrl_selftest_create: bad argument.");
    auto* cp = new igg::ControlPlane();
    std::string name;
    int rc = cp->create(me, nprocs, name);
    if (rc) { delete cp; return rc; }
    snprintf(name_out, 64, "%s", name.c_str());
    *handle = cp;
    return IGG_OK;
}

// `names`: nprocs entries of 64 bytes in rank order.  Runs `rounds` all-gathers whose payload encodes (rank, round)
// and verifies what comes back; then a barrier; then destroys the handle.
int igg_ctrl_selftest_run(void* handle, int me, int nprocs, const char* names, int rounds) {
    auto* cp = static_cast<igg::ControlPlane*>(handle);
    if (!cp || !names) return igg::fail(IGG_ERR_ARG, "igg_ctrl_selftest_run: bad argument.");
    std::vector<std::string> v;
    for (int r = 0; r < nprocs; ++r) v.emplace_back(names + (size_t)r * 64);
    int rc = cp->attach(v);
    for (int k = 0; k < rounds && !rc; ++k) {
        char mine[igg::kMailPayload];
        for (int i = 0; i < igg::kMailPayload; ++i) mine[i] = (char)((me * 31 + k * 7 + i) & 0xff);
        std::vector<char> all;
        rc = cp->allgather(mine, all);
        for (int r = 0; r < nprocs && !rc; ++r)
            for (int i = 0; i < igg::kMailPayload; ++i)
                if (all[(size_t)r * igg::kMailPayload + i] != (char)((r * 31 + k * 7 + i) & 0xff))
                    rc = igg::fail(IGG_ERR_IPC, "control plane self test: payload mismatch (rank %d round %d).", r, k);
    }
    if (!rc) rc = cp->barrier();
    delete cp;
    return rc;
}

}  // extern "C"
