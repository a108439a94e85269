// tools/host_path.cu — what the host side of sdrb_process_host can move: plain cudaMemcpyAsync bandwidth between
// pinned host memory and every visible GPU, one GPU at a time and all GPUs at once, host-to-device, device-to-host
// and both directions together.  Names the ceiling under the e2e numbers of bench.py.
//   nvcc -O2 -o tools/host_path.bin tools/host_path.cu ; tools/host_path.bin [MiB per copy = 256] [copies = 12]
// Prints one JSON object.  One host thread, one stream pair and its own pinned buffers per GPU.
#include <cuda_runtime.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

struct Gpu {
    int dev;
    void *h_in = nullptr, *h_out = nullptr, *d_a = nullptr, *d_b = nullptr;
    cudaStream_t s_in = nullptr, s_out = nullptr;
    char bus[32] = {0};
    int numa = -1;
};

static double now() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// mode: 1 = H2D, 2 = D2H, 3 = both at once.  Returns GB/s per direction summed over the GPUs in `set`.
static void run(std::vector<Gpu>& g, const std::vector<int>& set, int mode, size_t bytes, int copies, double* h2d,
                double* d2h) {
    std::atomic<int> ready{0};
    std::atomic<bool> go{false};
    std::vector<double> t((size_t)set.size());
    std::vector<std::thread> th;
    for (size_t k = 0; k < set.size(); k++)
        th.emplace_back([&, k] {
            Gpu& x = g[(size_t)set[k]];
            cudaSetDevice(x.dev);
            for (int w = 0; w < 2; w++) {   // warm-up
                if (mode & 1) cudaMemcpyAsync(x.d_a, x.h_in, bytes, cudaMemcpyHostToDevice, x.s_in);
                if (mode & 2) cudaMemcpyAsync(x.h_out, x.d_b, bytes, cudaMemcpyDeviceToHost, x.s_out);
            }
            cudaStreamSynchronize(x.s_in);
            cudaStreamSynchronize(x.s_out);
            ready++;
            while (!go) {}
            const double t0 = now();
            for (int c = 0; c < copies; c++) {
                if (mode & 1) cudaMemcpyAsync(x.d_a, x.h_in, bytes, cudaMemcpyHostToDevice, x.s_in);
                if (mode & 2) cudaMemcpyAsync(x.h_out, x.d_b, bytes, cudaMemcpyDeviceToHost, x.s_out);
            }
            cudaStreamSynchronize(x.s_in);
            cudaStreamSynchronize(x.s_out);
            t[k] = now() - t0;
        });
    while (ready < (int)set.size()) {}
    go = true;
    for (auto& x : th) x.join();
    double tmax = 0;
    for (double v : t) tmax = v > tmax ? v : tmax;
    const double gb = (double)bytes * copies * set.size() / 1e9;
    *h2d = (mode & 1) ? gb / tmax : 0.0;
    *d2h = (mode & 2) ? gb / tmax : 0.0;
}

int main(int argc, char** argv) {
    const size_t bytes = (size_t)(argc > 1 ? atoi(argv[1]) : 256) << 20;
    const int copies = argc > 2 ? atoi(argv[2]) : 12;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) { fprintf(stderr, "no CUDA device\n"); return 1; }
    std::vector<Gpu> g((size_t)n);
    for (int i = 0; i < n; i++) {
        Gpu& x = g[(size_t)i];
        x.dev = i;
        cudaSetDevice(i);
        cudaDeviceGetPCIBusId(x.bus, sizeof x.bus, i);
        std::string p = std::string("/sys/bus/pci/devices/") + x.bus + "/numa_node";
        for (auto& c : p) c = (char)tolower(c);
        if (FILE* f = fopen(p.c_str(), "r")) { if (fscanf(f, "%d", &x.numa) != 1) x.numa = -1; fclose(f); }
        if (cudaHostAlloc(&x.h_in, bytes, cudaHostAllocDefault) != cudaSuccess ||
            cudaHostAlloc(&x.h_out, bytes, cudaHostAllocDefault) != cudaSuccess ||
            cudaMalloc(&x.d_a, bytes) != cudaSuccess || cudaMalloc(&x.d_b, bytes) != cudaSuccess) {
            fprintf(stderr, "allocation failed on GPU %d\n", i);
            return 1;
        }
        memset(x.h_in, 1, bytes);
        memset(x.h_out, 0, bytes);
        cudaStreamCreateWithFlags(&x.s_in, cudaStreamNonBlocking);
        cudaStreamCreateWithFlags(&x.s_out, cudaStreamNonBlocking);
    }
    printf("{\"mib_per_copy\": %zu, \"copies\": %d, \"gpus\": %d, \"host_threads\": %u,\n \"per_gpu\": [", bytes >> 20, copies, n,
           std::thread::hardware_concurrency());
    for (int i = 0; i < n; i++) {
        double a, b, c, d, e, f;
        run(g, {i}, 1, bytes, copies, &a, &b);
        run(g, {i}, 2, bytes, copies, &c, &d);
        run(g, {i}, 3, bytes, copies, &e, &f);
        printf("%s\n  {\"gpu\": %d, \"pci\": \"%s\", \"numa_node\": %d, \"h2d_gbs\": %.2f, \"d2h_gbs\": %.2f, "
               "\"bidir_h2d_gbs\": %.2f, \"bidir_d2h_gbs\": %.2f}",
               i ? "," : "", i, g[(size_t)i].bus, g[(size_t)i].numa, a, d, e, f);
    }
    printf("],\n \"concurrent\": [");
    bool first = true;
    for (int k = 2; k <= n; k *= 2) {
        std::vector<int> set;
        for (int i = 0; i < k; i++) set.push_back(i);
        double a, b, c, d, e, f;
        run(g, set, 1, bytes, copies, &a, &b);
        run(g, set, 2, bytes, copies, &c, &d);
        run(g, set, 3, bytes, copies, &e, &f);
        printf("%s\n  {\"gpus\": %d, \"h2d_gbs\": %.2f, \"d2h_gbs\": %.2f, \"bidir_h2d_gbs\": %.2f, \"bidir_d2h_gbs\": %.2f, "
               "\"bidir_sum_gbs\": %.2f}", first ? "" : ",", k, a, d, e, f, e + f);
        first = false;
    }
    printf("]}\n");
    return 0;
}
