// rendezvous_test.cpp — TEST ONLY: the product's host-side rendezvous (csrc/rendezvous.h) with real
// processes and no GPU.  N forked ranks meet on an abstract-namespace socket, rank 0 shares a file
// descriptor with the others (here: a pipe's write end instead of a CUDA multicast handle), each
// rank proves its copy refers to the same object by writing its rank through it, and barriers
// order the phases.  Usage: rendezvous_test <nranks> [missing_rank]
#include <cstdio>
#include <cstdlib>
#include <string>
#include <sys/wait.h>
#include <unistd.h>
#include <vector>

#include "rendezvous.h"

static int run_rank(const std::string &name, int rank, int nranks, int pipe_w) {
    b200v::Rendezvous rv;
    if (rv.open(name.c_str(), rank, nranks) != b200v::kRvOk) return 10;
    if (!rv.barrier()) return 11;
    int fd = rank == 0 ? pipe_w : -1;
    if (!rv.share_fd(fd)) return 12;
    if (fd < 0) return 13;
    const char byte = (char)('0' + rank);
    if (::write(fd, &byte, 1) != 1) return 14;          // every rank writes through ITS descriptor
    if (!rv.barrier()) return 15;
    if (!rv.barrier()) return 16;                         // barriers are reusable

    // all-to-all: every rank contributes the write end of a pipe of its own and receives everyone
    // else's; each rank then writes its rank into every OTHER rank's pipe and must find exactly the
    // other ranks' bytes in its own.
    int mine[2];
    if (::pipe(mine) != 0) return 17;
    std::vector<int> all;
    if (!rv.exchange_fds(mine[1], all)) return 18;
    if ((int)all.size() != nranks || all[(size_t)rank] != mine[1]) return 19;
    for (int j = 0; j < nranks; ++j)
        if (j != rank && ::write(all[(size_t)j], &byte, 1) != 1) return 20;
    if (!rv.barrier()) return 21;
    std::vector<int> seen((size_t)nranks, 0);
    for (int j = 0; j < nranks - 1; ++j) {
        char got = 0;
        if (::read(mine[0], &got, 1) != 1) return 22;
        seen[(size_t)(got - '0')]++;
    }
    for (int j = 0; j < nranks; ++j)
        if (seen[(size_t)j] != (j == rank ? 0 : 1)) return 23;
    for (int j = 0; j < nranks; ++j)
        if (j != rank) ::close(all[(size_t)j]);
    rv.close_all();
    return 0;
}

int main(int argc, char **argv) {
    const int nranks = argc > 1 ? std::atoi(argv[1]) : 3;
    const int missing = argc > 2 ? std::atoi(argv[2]) : -1;
    const std::string name = "b200v-rv-test-" + std::to_string((long)::getpid());
    int p[2];
    if (::pipe(p) != 0) return 2;
    std::vector<pid_t> pids;
    for (int r = 0; r < nranks; ++r) {
        if (r == missing) continue;
        pid_t pid = ::fork();
        if (pid == 0) {
            ::close(p[0]);
            const int w = r == 0 ? p[1] : -1;
            if (r != 0) ::close(p[1]);                    // only rank 0 starts out holding the descriptor
            ::_exit(run_rank(name, r, nranks, w));
        }
        pids.push_back(pid);
    }
    ::close(p[1]);
    int failures = 0;
    for (pid_t pid : pids) {
        int st = 0;
        ::waitpid(pid, &st, 0);
        if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) {
            std::printf("rank process exit %d\n", WIFEXITED(st) ? WEXITSTATUS(st) : -1);
            ++failures;
        }
    }
    char buf[64];
    ssize_t got = ::read(p[0], buf, sizeof buf);
    std::vector<int> seen((size_t)nranks, 0);
    for (ssize_t i = 0; i < got; ++i) seen[(size_t)(buf[i] - '0')]++;
    bool all = got == nranks;
    for (int r = 0; r < nranks; ++r) all = all && seen[(size_t)r] == 1;
    std::printf("%s failures=%d bytes=%zd\n", (failures == 0 && all) ? "RENDEZVOUS_OK" : "RENDEZVOUS_BAD", failures, got);
    return (failures == 0 && all) ? 0 : 1;
}
