// rendezvous.h — host-side meeting point of the ranks of one node, one process per GPU.
//
// The NVLink-multicast probe needs every process to hold the SAME multicast object.  Rank 0
// creates it, exports it as a POSIX file descriptor and hands that descriptor to the other ranks
// over a Unix-domain socket (SCM_RIGHTS); the same sockets carry the few host-side barriers the
// set-up needs.  The socket lives in the abstract namespace under the name the caller passes, so
// nothing is left on disk.  Every wait is bounded (kRvTimeoutMs): a missing rank makes the call
// fail, never hang.  Plain POSIX, no CUDA: tests/host/rendezvous_test.cpp exercises it on the CPU.
#pragma once

#include <algorithm>
#include <cstddef>
#include <cstdlib>
#include <cstring>
#include <poll.h>
#include <sys/socket.h>
#include <sys/uio.h>
#include <sys/un.h>
#include <unistd.h>
#include <vector>

namespace b200v {

constexpr int kRvOk = 0, kRvFailed = -10;     // kRvFailed == B200V_E_TIMEOUT

// 60 s by default; B200V_RV_TIMEOUT_MS shortens it (the CPU test of the missing-rank path does).
inline int rv_timeout_ms() {
    static const int ms = [] {
        const char *e = std::getenv("B200V_RV_TIMEOUT_MS");
        const int v = e && *e ? std::atoi(e) : 0;
        return v > 0 ? v : 60000;
    }();
    return ms;
}
#define kRvTimeoutMs (::b200v::rv_timeout_ms())

struct Rendezvous {
    int rank = -1, nranks = 0;
    int listen_fd = -1;
    int root_fd = -1;              // ranks > 0: connection to rank 0
    std::vector<int> peer_fd;      // rank 0: connection of each rank (index = rank; [0] unused)
    bool is_open = false;

    static bool wait_readable(int fd, int timeout_ms) {
        pollfd p{fd, POLLIN, 0};
        return ::poll(&p, 1, timeout_ms) == 1 && (p.revents & POLLIN);
    }
    static bool send_all(int fd, const void *buf, size_t n) {
        const char *p = (const char *)buf;
        while (n) {
            ssize_t w = ::send(fd, p, n, MSG_NOSIGNAL);
            if (w <= 0) return false;
            p += w;
            n -= (size_t)w;
        }
        return true;
    }
    static bool recv_all(int fd, void *buf, size_t n) {
        char *p = (char *)buf;
        while (n) {
            if (!wait_readable(fd, kRvTimeoutMs)) return false;
            ssize_t r = ::recv(fd, p, n, 0);
            if (r <= 0) return false;
            p += r;
            n -= (size_t)r;
        }
        return true;
    }
    static socklen_t make_addr(const char *name, sockaddr_un &a) {
        std::memset(&a, 0, sizeof a);
        a.sun_family = AF_UNIX;
        const size_t len = std::min(std::strlen(name), sizeof(a.sun_path) - 2);
        std::memcpy(a.sun_path + 1, name, len);            // leading NUL = abstract namespace
        return (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + len);
    }

    int open(const char *name, int rank_, int nranks_) {
        rank = rank_;
        nranks = nranks_;
        sockaddr_un addr;
        const socklen_t alen = make_addr(name, addr);
        if (rank == 0) {
            listen_fd = ::socket(AF_UNIX, SOCK_STREAM, 0);
            if (listen_fd < 0 || ::bind(listen_fd, (sockaddr *)&addr, alen) != 0 || ::listen(listen_fd, nranks) != 0)
                return kRvFailed;
            peer_fd.assign((size_t)nranks, -1);
            for (int k = 1; k < nranks; ++k) {
                if (!wait_readable(listen_fd, kRvTimeoutMs)) return kRvFailed;
                int c = ::accept(listen_fd, nullptr, nullptr);
                int who = -1;
                if (c < 0 || !recv_all(c, &who, sizeof who) || who < 1 || who >= nranks || peer_fd[(size_t)who] >= 0) {
                    if (c >= 0) ::close(c);
                    return kRvFailed;
                }
                peer_fd[(size_t)who] = c;
            }
        } else {
            for (int waited = 0;; waited += 20) {              // rank 0 may not be listening yet
                root_fd = ::socket(AF_UNIX, SOCK_STREAM, 0);
                if (root_fd >= 0 && ::connect(root_fd, (sockaddr *)&addr, alen) == 0) break;
                if (root_fd >= 0) ::close(root_fd);
                root_fd = -1;
                if (waited >= kRvTimeoutMs) return kRvFailed;
                ::usleep(20000);
            }
            if (!send_all(root_fd, &rank, sizeof rank)) return kRvFailed;
        }
        is_open = true;
        return kRvOk;
    }

    // All ranks meet (host side).  Returns false if a rank is missing.
    bool barrier() {
        if (!is_open) return false;
        char b = 'b';
        if (rank == 0) {
            for (int k = 1; k < nranks; ++k)
                if (!recv_all(peer_fd[(size_t)k], &b, 1)) return false;
            for (int k = 1; k < nranks; ++k)
                if (!send_all(peer_fd[(size_t)k], &b, 1)) return false;
            return true;
        }
        return send_all(root_fd, &b, 1) && recv_all(root_fd, &b, 1);
    }

    // One descriptor over one socket (SCM_RIGHTS with a one-byte payload).
    static bool send_fd(int sock, int fd) {
        char tag = 'f';
        iovec iov{&tag, 1};
        alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))];
        std::memset(ctrl, 0, sizeof ctrl);
        msghdr msg;
        std::memset(&msg, 0, sizeof msg);
        msg.msg_iov = &iov;
        msg.msg_iovlen = 1;
        msg.msg_control = ctrl;
        msg.msg_controllen = sizeof ctrl;
        cmsghdr *c = CMSG_FIRSTHDR(&msg);
        c->cmsg_level = SOL_SOCKET;
        c->cmsg_type = SCM_RIGHTS;
        c->cmsg_len = CMSG_LEN(sizeof(int));
        std::memcpy(CMSG_DATA(c), &fd, sizeof(int));
        return ::sendmsg(sock, &msg, MSG_NOSIGNAL) == 1;
    }
    static bool recv_fd(int sock, int &fd) {
        char tag = 0;
        iovec iov{&tag, 1};
        alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))];
        msghdr msg;
        std::memset(&msg, 0, sizeof msg);
        msg.msg_iov = &iov;
        msg.msg_iovlen = 1;
        msg.msg_control = ctrl;
        msg.msg_controllen = sizeof ctrl;
        if (!wait_readable(sock, kRvTimeoutMs) || ::recvmsg(sock, &msg, 0) != 1) return false;
        cmsghdr *c = CMSG_FIRSTHDR(&msg);
        if (!c || c->cmsg_level != SOL_SOCKET || c->cmsg_type != SCM_RIGHTS) return false;
        std::memcpy(&fd, CMSG_DATA(c), sizeof(int));
        return fd >= 0;
    }

    // Rank 0 gives `fd` to every other rank; the others receive their own copy of it.
    bool share_fd(int &fd) {
        if (!is_open) return false;
        if (rank == 0) {
            for (int k = 1; k < nranks; ++k)
                if (!send_fd(peer_fd[(size_t)k], fd)) return false;
            return true;
        }
        return recv_fd(root_fd, fd);
    }

    // Every rank contributes one descriptor and ends up holding a copy of every rank's, in rank
    // order (all[rank] is the caller's own `mine`, not a copy).  Rank 0 relays.  The caller closes
    // the copies it received.
    bool exchange_fds(int mine, std::vector<int> &all) {
        if (!is_open) return false;
        all.assign((size_t)nranks, -1);
        all[(size_t)rank] = mine;
        if (rank == 0) {
            for (int k = 1; k < nranks; ++k)
                if (!recv_fd(peer_fd[(size_t)k], all[(size_t)k])) return false;
            for (int k = 1; k < nranks; ++k)
                for (int j = 0; j < nranks; ++j)
                    if (j != k && !send_fd(peer_fd[(size_t)k], all[(size_t)j])) return false;
            return true;
        }
        if (!send_fd(root_fd, mine)) return false;
        for (int j = 0; j < nranks; ++j)
            if (j != rank && !recv_fd(root_fd, all[(size_t)j])) return false;
        return true;
    }

    void close_all() {
        for (int f : peer_fd)
            if (f >= 0) ::close(f);
        peer_fd.clear();
        if (root_fd >= 0) ::close(root_fd);
        if (listen_fd >= 0) ::close(listen_fd);
        root_fd = listen_fd = -1;
        is_open = false;
    }
};

}  // namespace b200v
