// lg_pair_bench — batched multi-GPU pair-matching bench harness (the north star's "new batched multi-GPU
// matching bench harness beside demo/"; the reference's own driver is demo/demo_mono.cpp, one pair at a time).
//
//   lg_pair_bench --blob weights.blob [--layers 9] [--gpus 1] [--kpts 2048] [--pairs 64] [--sub 32] [--steps 5]
//
// One PROCESS per GPU (forked before any CUDA call), each with its own engine and stream; image pairs are
// independent, so pairs are sharded p -> GPU p mod G with no collective; every worker sends its per-pair match
// counts and its elapsed time to the parent over a pipe, and the parent prints ONE JSON line (job time = the
// slowest worker's).  Features are synthetic SuperPoint-shaped: normalised keypoints,
// L2-normalised 256-d FP16 descriptors, in pinned host memory; the timed region covers H2D + forward + D2H
// through lg_matcher_submit_host / lg_matcher_sync (inputs of sub-batch i+1 staged under forward i).  Pure C++ over the C-ABI (include/lg_matcher.h): no Python, torch or TensorRT.
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <sched.h>
#include <sys/wait.h>
#include <unistd.h>

#include <cctype>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <vector>

#include "../../../include/lg_attn.h"
#include "../../../include/lg_matcher.h"

struct Args {
    std::string blob;
    int layers = 9, gpus = 1, kpts = 2048, pairs = 64, sub = 32, steps = 5;
};

struct WorkerResult {
    double seconds;
    long long pairs_done;
    long long mutual_matches;
    int rc;
};

// Run this worker on the CPUs local to its GPU (/sys/bus/pci/devices/<pci>/local_cpulist), so that the pinned
// host buffers it first-touches next sit on the GPU's own NUMA node.  Best effort; returns the number of CPUs
// bound to, 0 when nothing was changed.
static int bind_to_gpu_local_cpus(int device) {
    char bus[32] = {0};
    if (cudaDeviceGetPCIBusId(bus, sizeof(bus), device) != cudaSuccess) { (void)cudaGetLastError(); return 0; }
    for (char* c = bus; *c; ++c) *c = static_cast<char>(tolower(*c));
    const std::string path = std::string("/sys/bus/pci/devices/") + bus + "/local_cpulist";
    FILE* f = fopen(path.c_str(), "r");
    if (f == nullptr) return 0;
    char line[4096] = {0};
    const bool ok = fgets(line, sizeof(line), f) != nullptr;
    fclose(f);
    if (!ok) return 0;
    cpu_set_t allowed, target;
    if (sched_getaffinity(0, sizeof(allowed), &allowed) != 0) return 0;
    CPU_ZERO(&target);
    int n = 0;
    for (char* tok = strtok(line, ",\n"); tok != nullptr; tok = strtok(nullptr, ",\n")) {
        int lo = 0, hi = 0;
        const int got = sscanf(tok, "%d-%d", &lo, &hi);
        if (got < 1) continue;
        if (got == 1) hi = lo;
        for (int c = lo; c <= hi && c < CPU_SETSIZE; ++c)
            if (c >= 0 && CPU_ISSET(c, &allowed)) { CPU_SET(c, &target); ++n; }
    }
    if (n == 0 || sched_setaffinity(0, sizeof(target), &target) != 0) return 0;
    return n;
}

static int worker(const Args& a, int rank, const std::vector<char>& blob, WorkerResult* res) {
    res->rc = 0;
    lg_matcher* eng = nullptr;
    int rc = lg_matcher_create(&eng, rank, a.layers, blob.data(), blob.size(), 0.0f);
    if (rc != 0) { fprintf(stderr, "rank %d: lg_matcher_create: %s\n", rank, lg_attn_error_string(rc)); return rc; }
    if (getenv("LG_BENCH_NO_BIND") == nullptr) (void)bind_to_gpu_local_cpus(rank);
    const int n = a.kpts, sub = a.sub < a.pairs ? a.sub : a.pairs;
    // pinned host features for this rank's pairs (global pair id p = rank + i * gpus seeds the generator)
    float *k0, *k1, *ms;
    __half *d0, *d1;
    int* m0;
    const size_t nk = static_cast<size_t>(a.pairs) * n * 2, nd = static_cast<size_t>(a.pairs) * n * 256, nm = static_cast<size_t>(a.pairs) * n;
    {
        cudaError_t me = cudaMallocHost(&k0, nk * 4);
        if (me == cudaSuccess) me = cudaMallocHost(&k1, nk * 4);
        if (me == cudaSuccess) me = cudaMallocHost(&d0, nd * 2);
        if (me == cudaSuccess) me = cudaMallocHost(&d1, nd * 2);
        if (me == cudaSuccess) me = cudaMallocHost(&m0, nm * 4);
        if (me == cudaSuccess) me = cudaMallocHost(&ms, nm * 4);
        if (me != cudaSuccess) {
            fprintf(stderr, "rank %d: cudaMallocHost: %s\n", rank, cudaGetErrorString(me));
            lg_matcher_destroy(eng);
            return static_cast<int>(me);
        }
    }
    for (int i = 0; i < a.pairs; ++i) {
        std::mt19937 gen(1000u + static_cast<unsigned>(rank + i * a.gpus));
        std::uniform_real_distribution<float> u(-1.f, 1.f);
        std::normal_distribution<float> g(0.f, 1.f);
        for (int img = 0; img < 2; ++img) {
            float* k = (img ? k1 : k0) + static_cast<size_t>(i) * n * 2;
            __half* d = (img ? d1 : d0) + static_cast<size_t>(i) * n * 256;
            for (int r = 0; r < n; ++r) {
                k[2 * r] = u(gen);
                k[2 * r + 1] = 0.75f * u(gen);
                float row[256], ss = 0.f;
                for (int c = 0; c < 256; ++c) { row[c] = g(gen); ss += row[c] * row[c]; }
                const float inv = 1.0f / std::sqrt(ss);
                for (int c = 0; c < 256; ++c) d[static_cast<size_t>(r) * 256 + c] = __float2half(row[c] * inv);
            }
        }
    }
    if ((rc = lg_matcher_plan(eng, sub, n, n)) != 0) { fprintf(stderr, "rank %d: plan: %s\n", rank, lg_attn_error_string(rc)); return rc; }
    auto step = [&]() {
        for (int s = 0; s < a.pairs && rc == 0; s += sub) {
            const int b = (a.pairs - s) < sub ? (a.pairs - s) : sub;
            rc = lg_matcher_submit_host(eng, k0 + static_cast<size_t>(s) * n * 2, k1 + static_cast<size_t>(s) * n * 2,
                                       d0 + static_cast<size_t>(s) * n * 256, d1 + static_cast<size_t>(s) * n * 256, b, n, n,
                                       m0 + static_cast<size_t>(s) * n, ms + static_cast<size_t>(s) * n);
        }
        if (rc == 0) rc = lg_matcher_sync(eng);   // one step = all of this rank's pairs matched and back on the host
    };
    step();  // warm-up (a short last sub-batch gets its own plan here; afterwards the two plans are swapped, not rebuilt)
    step();
    const auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < a.steps && rc == 0; ++i) step();
    const auto t1 = std::chrono::steady_clock::now();
    res->seconds = std::chrono::duration<double>(t1 - t0).count();
    res->pairs_done = static_cast<long long>(a.pairs) * a.steps;
    long long mutual = 0;
    for (size_t i = 0; i < nm; ++i) mutual += m0[i] >= 0;
    res->mutual_matches = mutual;
    res->rc = rc;
    lg_matcher_destroy(eng);
    return rc;
}

int main(int argc, char** argv) {
    Args a;
    for (int i = 1; i + 1 < argc; i += 2) {
        const std::string k = argv[i], v = argv[i + 1];
        if (k == "--blob") a.blob = v;
        else if (k == "--layers") a.layers = atoi(v.c_str());
        else if (k == "--gpus") a.gpus = atoi(v.c_str());
        else if (k == "--kpts") a.kpts = atoi(v.c_str());
        else if (k == "--pairs") a.pairs = atoi(v.c_str());
        else if (k == "--sub") a.sub = atoi(v.c_str());
        else if (k == "--steps") a.steps = atoi(v.c_str());
        else { fprintf(stderr, "unknown option %s\n", k.c_str()); return 2; }
    }
    if (a.blob.empty() || a.gpus < 1 || a.kpts < 1 || a.pairs < 1 || a.sub < 1 || a.steps < 1) {
        fprintf(stderr, "usage: lg_pair_bench --blob FILE [--layers 9] [--gpus 1] [--kpts 2048] [--pairs 64] [--sub 32] [--steps 5]\n");
        return 2;
    }
    FILE* f = fopen(a.blob.c_str(), "rb");
    if (!f) { perror("blob"); return 2; }
    std::vector<char> blob(lg_matcher_blob_bytes(a.layers));
    const size_t got = fread(blob.data(), 1, blob.size(), f);
    fclose(f);
    if (got != blob.size()) { fprintf(stderr, "blob: expected %zu bytes, read %zu\n", blob.size(), got); return 2; }

    std::vector<int> fds(a.gpus);
    std::vector<pid_t> pids(a.gpus);
    for (int r = 0; r < a.gpus; ++r) {   // fork BEFORE any CUDA call: one process (and context) per GPU
        int p[2];
        if (pipe(p) != 0) { perror("pipe"); return 1; }
        const pid_t pid = fork();
        if (pid == 0) {
            close(p[0]);
            WorkerResult res{};
            const int rc = worker(a, r, blob, &res);
            res.rc = rc;
            (void)!write(p[1], &res, sizeof(res));
            _exit(rc == 0 ? 0 : 1);
        }
        close(p[1]);
        fds[r] = p[0];
        pids[r] = pid;
    }
    double worst = 0;
    long long pairs = 0, mutual = 0;
    int bad = 0;
    for (int r = 0; r < a.gpus; ++r) {   // gather to the host of the parent
        WorkerResult res{};
        if (read(fds[r], &res, sizeof(res)) != static_cast<ssize_t>(sizeof(res)) || res.rc != 0) ++bad;
        int st = 0;
        waitpid(pids[r], &st, 0);
        worst = res.seconds > worst ? res.seconds : worst;
        pairs += res.pairs_done;
        mutual += res.mutual_matches;
    }
    if (bad) { fprintf(stderr, "%d worker(s) failed\n", bad); return 1; }
    printf("{\"metric\": \"LightGlue pairs/s (native harness, host buffers)\", \"value\": %.1f, \"unit\": \"pairs/s\", \"n_gpus\": %d, "
           "\"kpts\": %d, \"pairs_per_gpu_per_step\": %d, \"pairs_per_forward\": %d, \"steps\": %d, \"seconds\": %.4f, "
           "\"mutual_nn_per_pair\": %.1f, \"sharding\": \"pair p -> GPU p mod G, one process per GPU, no collective\"}\n",
           pairs / worst, a.gpus, a.kpts, a.pairs, a.sub, a.steps, worst, static_cast<double>(mutual) / (a.pairs * a.gpus));
    return 0;
}
