// h2d_matrix.cu -- what this box can move from page-locked host memory to its GPUs, with no engine and no Python in
// the way: the T1 roof per GPU count ("at T1 the binding roof is PCIe Gen5 x16 H2D").
//
//   nvcc -O2 -o tools/h2d_matrix tools/h2d_matrix.cu -lpthread && tools/h2d_matrix [seconds per point] > profiles/h2d_matrix_r02.jsonl
//
// One host thread per GPU, each with its own pinned buffer and stream, all released together; every point runs for
// `seconds` (default 2) of back-to-back 96 MiB cudaMemcpyAsync H2D. Points: each GPU alone, every GPU pair, the first
// 1/2/4/8 GPUs together, and the same with the threads pinned to the CPUs of the GPU's NUMA node when sysfs tells
// (pinned = 1) and with write-combined pinned memory (wc = 1). One JSON line per point.
#include <cuda_runtime.h>
#include <pthread.h>
#include <sched.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

static const size_t kBytes = 96ull << 20;

struct Point { std::vector<int> gpus; bool affinity; bool wc; };

static std::vector<int> cpus_of_gpu(int dev) {  // CPUs local to the GPU's PCI device, from sysfs; empty if unknown
    char bus[32];
    std::vector<int> out;
    if (cudaDeviceGetPCIBusId(bus, sizeof bus, dev) != cudaSuccess) return out;
    for (char* p = bus; *p; p++) *p = (char)tolower(*p);
    std::string path = std::string("/sys/bus/pci/devices/") + bus + "/local_cpulist";
    FILE* f = fopen(path.c_str(), "r");
    if (!f) return out;
    char line[512] = {0};
    if (fgets(line, sizeof line, f)) {
        for (char* tok = strtok(line, ",\n"); tok; tok = strtok(nullptr, ",\n")) {
            int a, b;
            if (sscanf(tok, "%d-%d", &a, &b) == 2) for (int c = a; c <= b; c++) out.push_back(c);
            else if (sscanf(tok, "%d", &a) == 1) out.push_back(a);
        }
    }
    fclose(f);
    return out;
}

static double run_point(const Point& pt, double seconds, std::vector<double>* per_gpu) {
    const int n = (int)pt.gpus.size();
    std::atomic<int> ready{0};
    std::atomic<bool> go{false}, stop{false};
    std::vector<double> gbs((size_t)n, 0.0);
    std::vector<std::thread> th;
    for (int i = 0; i < n; i++)
        th.emplace_back([&, i] {
            const int dev = pt.gpus[(size_t)i];
            if (pt.affinity) {
                std::vector<int> cpus = cpus_of_gpu(dev);
                if (!cpus.empty()) {
                    cpu_set_t set;
                    CPU_ZERO(&set);
                    for (int c : cpus) CPU_SET(c, &set);
                    pthread_setaffinity_np(pthread_self(), sizeof set, &set);
                }
            }
            cudaSetDevice(dev);
            void *h = nullptr, *d = nullptr;
            cudaHostAlloc(&h, kBytes, pt.wc ? cudaHostAllocWriteCombined : cudaHostAllocDefault);
            cudaMalloc(&d, kBytes);
            memset(h, 1, kBytes);  // first touch on this thread
            cudaStream_t st;
            cudaStreamCreate(&st);
            cudaEvent_t e0, e1;
            cudaEventCreate(&e0);
            cudaEventCreate(&e1);
            cudaMemcpyAsync(d, h, kBytes, cudaMemcpyHostToDevice, st);
            cudaStreamSynchronize(st);
            ready++;
            while (!go) std::this_thread::yield();
            size_t moved = 0;
            cudaEventRecord(e0, st);
            while (!stop) {
                for (int k = 0; k < 4; k++) cudaMemcpyAsync(d, h, kBytes, cudaMemcpyHostToDevice, st);
                cudaStreamSynchronize(st);
                moved += 4 * kBytes;
            }
            cudaEventRecord(e1, st);
            cudaEventSynchronize(e1);
            float ms = 0;
            cudaEventElapsedTime(&ms, e0, e1);
            gbs[(size_t)i] = (double)moved / (ms * 1e-3) * 1e-9;
            cudaFreeHost(h);
            cudaFree(d);
            cudaStreamDestroy(st);
        });
    while (ready < n) std::this_thread::yield();
    go = true;
    std::this_thread::sleep_for(std::chrono::duration<double>(seconds));
    stop = true;
    for (auto& t : th) t.join();
    double sum = 0;
    for (double g : gbs) sum += g;
    *per_gpu = gbs;
    return sum;
}

int main(int argc, char** argv) {
    const double seconds = argc > 1 ? atof(argv[1]) : 2.0;
    int ngpu = 0;
    cudaGetDeviceCount(&ngpu);
    if (ngpu <= 0) { fprintf(stderr, "no GPU\n"); return 1; }
    cudaDeviceProp pr;
    cudaGetDeviceProperties(&pr, 0);
    printf("{\"device\": \"%s\", \"gpus\": %d, \"host_cpus\": %u, \"bytes_per_copy\": %zu, \"seconds_per_point\": %.1f}\n", pr.name, ngpu,
           std::thread::hardware_concurrency(), kBytes, seconds);
    for (int g = 0; g < ngpu; g++) {
        std::vector<int> c = cpus_of_gpu(g);
        printf("{\"gpu\": %d, \"local_cpus\": %zu, \"first_cpu\": %d, \"last_cpu\": %d}\n", g, c.size(), c.empty() ? -1 : c.front(), c.empty() ? -1 : c.back());
    }
    std::vector<Point> pts;
    for (int g = 0; g < ngpu; g++) pts.push_back(Point{{g}, false, false});
    for (int a = 0; a < ngpu; a++)
        for (int b = a + 1; b < ngpu; b++) pts.push_back(Point{{a, b}, false, false});
    for (int k = 1; k <= ngpu; k *= 2) {
        std::vector<int> v;
        for (int g = 0; g < k; g++) v.push_back(g);
        if (k > 2 || ngpu == 1) pts.push_back(Point{v, false, false});
        pts.push_back(Point{v, true, false});
        pts.push_back(Point{v, false, true});
    }
    if (ngpu >= 8) pts.push_back(Point{{0, 2, 4, 6}, false, false}), pts.push_back(Point{{4, 5, 6, 7}, false, false});
    for (const Point& pt : pts) {
        std::vector<double> per;
        const double sum = run_point(pt, seconds, &per);
        printf("{\"gpus\": [");
        for (size_t i = 0; i < pt.gpus.size(); i++) printf("%s%d", i ? ", " : "", pt.gpus[i]);
        printf("], \"n\": %zu, \"cpu_affinity\": %d, \"write_combined\": %d, \"h2d_gbs_total\": %.2f, \"per_gpu\": [", pt.gpus.size(), pt.affinity ? 1 : 0,
               pt.wc ? 1 : 0, sum);
        for (size_t i = 0; i < per.size(); i++) printf("%s%.2f", i ? ", " : "", per[i]);
        printf("]}\n");
        fflush(stdout);
    }
    return 0;
}
