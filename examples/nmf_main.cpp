// nmf_main.cpp — the reference's CLI flow (cuda/nmf.cu:30-51) on top of libnmf_b200.so:
//   reads ../X.bin ../H.bin ../W.bin, runs update_div, writes ../Wout.bin ../Hout.bin.
// Optional arguments (the reference hard-codes all of these, nmf.cu:9-11,37-45):
//   nmf [dir=..] [max_iter=200] [converge_thresh=0] [verbose=0] [gpus=1]
// gpus > 1: update_div's signature has no GPU count (README.md:39-47), the library reads it from
// NMF_NUM_GPUS; this argument sets that variable (column-sharded run on that many GPUs of this box).
#include <cstdio>
#include <cstdlib>
#include <string>

#include "../include/nmf.h"

int main(int argc, char** argv) {
    const std::string dir = argc > 1 ? argv[1] : "..";
    const int max_iter = argc > 2 ? std::atoi(argv[2]) : 200;        // MAX_ITER, nmf.cu:10
    const float thresh = argc > 3 ? (float)std::atof(argv[3]) : 0.f; // CONVERGE_THRESH, nmf.cu:11
    const int verbose = argc > 4 ? std::atoi(argv[4]) : 0;
    if (argc > 5 && std::atoi(argv[5]) > 0) setenv("NMF_NUM_GPUS", argv[5], 1);
    matrix X, H, W;
    read_matrix(&X, (dir + "/X.bin").c_str());
    read_matrix(&H, (dir + "/H.bin").c_str());
    read_matrix(&W, (dir + "/W.bin").c_str());
    double t[10];
    update_div(W, H, X, thresh, max_iter, verbose ? t : nullptr, verbose);
    if (verbose)
        std::printf("seconds: WH-gemm %.4f sums %.4f update-gemms %.4f W-update %.4f cost %.4f comm %.4f copy/setup %.4f total %.4f\n",
                    t[0], t[2], t[3], t[5], t[6], t[7], t[8], t[9]);
    write_matrix(W, (dir + "/Wout.bin").c_str());
    write_matrix(H, (dir + "/Hout.bin").c_str());
    destroy_matrix(&X); destroy_matrix(&H); destroy_matrix(&W);
    return 0;
}
