// RankSync -- status rendezvous of the ranks of one multi-process abub3hs run (no counterpart in the reference, which is a
// single process).
#pragma once
#include <stdlib.h>
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

/* Multi-rank runs: the ranks agree on a status word before every phase that ends in a collective, so that a rank which
 * failed (no readable frame in its range, a CUDA error, a failed training) cannot leave its peers blocked inside
 * ncclCommInitRank / ncclAllReduce.  File rendezvous next to the output file: rank r publishes
 * <base>.<phase>.<r> = "<start time> <status>" (written whole, then renamed) and waits -- ABUB_RANK_TIMEOUT seconds,
 * default 300 -- for the other ranks' files of the same phase.  Files older than this rank's start minus two minutes are
 * leftovers of an earlier run and are ignored.  Returns the minimum status over the ranks, or -1 when a rank did not
 * show up in time.  A rank removes its own file of a phase once the next phase (or the final barrier) has passed. */
struct RankSync {
    std::string base;
    int rank = 0, world = 1;
    long long t_start = 0;
    std::vector<std::string> mine;
    double timeout() const { const char* e = getenv("ABUB_RANK_TIMEOUT"); const double t = e ? atof(e) : 0; return t > 0 ? t : 300.0; }
    int agree(const char* phase, int status) {
        if (world <= 1) return status;
        const std::string me = base + "." + phase + "." + std::to_string(rank);
        {
            std::ofstream f(me + ".tmp");
            f << t_start << " " << status << "\n";
        }
        rename((me + ".tmp").c_str(), me.c_str());
        mine.push_back(me);
        const auto t0 = std::chrono::steady_clock::now();
        int lo = status;
        for (int r = 0; r < world; r++) {
            if (r == rank) continue;
            const std::string fn = base + "." + phase + "." + std::to_string(r);
            for (;;) {
                long long ts = 0;
                int st = 0;
                std::ifstream f(fn);
                if (f && (f >> ts >> st) && ts >= t_start - 120) { lo = std::min(lo, st); break; }
                if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > timeout()) {
                    std::cout << "rank " << rank << ": rank " << r << " did not reach phase '" << phase << "' in time\n";
                    return -1;
                }
                usleep(20000);
            }
        }
        return lo;
    }
    void forget(size_t keep_last) {
        while (mine.size() > keep_last) { unlink(mine.front().c_str()); mine.erase(mine.begin()); }
    }
};

