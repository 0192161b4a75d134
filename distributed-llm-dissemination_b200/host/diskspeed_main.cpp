// `diskspeed` — the reference's disk-speed probe (diskspeed/main.go: -path <file>; prints file size, time to load and
// MiB/s of reading the file into memory) with the device store as the destination: the file is attached as a node's
// disk copy and loaded into that node's HBM through the staging ring of the disk tier (dissem_send_file), so the figure
// is what a DiskLayer source can deliver on this box, checksum pass included.
#include <errno.h>
#include <sys/stat.h>

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../../include/dissem.h"

int main(int argc, char **argv) {
    std::string path;
    long long threads = 0;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        if ((a == "-path" || a == "--path") && i + 1 < argc) path = argv[++i];
        else if (a == "--threads" && i + 1 < argc) threads = atoll(argv[++i]);
    }
    if (path.empty()) {                                                  // diskspeed/main.go:23-26
        printf("usage: -path [file path]\n");
        return 0;
    }
    struct stat st;
    if (stat(path.c_str(), &st) != 0) {
        fprintf(stderr, "failed to get fileInfo: %s\n", strerror(errno));
        return 1;
    }
    const int64_t size = (int64_t)st.st_size;
    if (dissem_device_count() <= 0) {
        fprintf(stderr, "no CUDA device: this tool loads into HBM and has no CPU path\n");
        return 1;
    }
    const int dev = 0;
    dissem_ctx *ctx = nullptr;
    auto fail = [&](const char *what) {
        fprintf(stderr, "failed to load: %s: %s\n", what, dissem_last_error());
        if (ctx) dissem_close(ctx);
        return 1;
    };
    if (dissem_init(1, &dev, &ctx)) return fail("dissem_init");
    if (threads > 0 && dissem_set_option(ctx, "file_threads", threads)) return fail("file_threads");
    if (dissem_layer_attach_file(ctx, 0, 0, path.c_str(), 0, size)) return fail("dissem_layer_attach_file");
    if (dissem_layer_alloc(ctx, 0, 0, size)) return fail("dissem_layer_alloc");
    const auto t0 = std::chrono::steady_clock::now();                    // diskspeed/main.go:33: the clock starts before the read
    dissem_ticket ticket = 0;
    if (dissem_send_file(ctx, 0, 0, 0, 0, size, size, &ticket)) return fail("dissem_send_file");
    if (dissem_wait(ctx, ticket, nullptr)) return fail("dissem_wait");
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    int64_t landed = 0;
    if (dissem_layer_received(ctx, 0, 0, &landed) || landed != size) return fail("dissem_layer_received");
    printf("File size: %lld\n", (long long)size);                        // diskspeed/main.go:42-44
    printf("Time to load: %.6fms\n", secs * 1e3);
    printf("Throughput: %.2f MiB/s\n", (double)size / secs / std::pow(2.0, 20));
    dissem_close(ctx);
    return 0;
}
