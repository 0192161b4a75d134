// `dissem` — the reference's CLI surface (cmd/main.go:15-21: -id -f -s -m -l -c -v) over the device engine.
// The reference starts one process per node; here every node of the config is a GPU (or a slice of one)
// of this box and all of them run inside this process, so -id only selects whose view is printed.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>

#include "../../include/dissem_host.h"

int main(int argc, char **argv) {
    int id = -1, mode = -1, verbose = 0, batched = 0, gpus = 1, pace = 0, dump_tables = 0;
    std::string file, storage;
    bool save_disk = false;
    long long time_scale = 1000000;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto next = [&]() -> const char * { return i + 1 < argc ? argv[++i] : ""; };
        if (a == "-id") id = atoi(next());
        else if (a == "-f") file = next();
        else if (a == "-s") { storage = next(); save_disk = !storage.empty(); }   // cmd/main.go:93-96: layers live in files under this path
        else if (a == "-m") mode = atoi(next());
        else if (a == "-l") { printf("layer set up\n"); return 0; }   // cmd/main.go:108-111: layers are synthesised in HBM at start
        else if (a == "-c") { fprintf(stderr, "external clients are outside the device transport's scope\n"); return 2; }
        else if (a == "-v") verbose = 1;
        else if (a == "--batched") batched = 1;
        else if (a == "--pace") pace = 1;                             // honour LimitRate as writeWithLimit does (transport.go:407-424)
        else if (a == "--dump-tables") dump_tables = 1;              // with -v: the report carries every delivered layer's table and edge bytes
        else if (a == "--gpus") gpus = atoi(next());
        else if (a == "--time-scale") time_scale = atoll(next());
    }
    if (id < 0 || file.empty()) {                                      // cmd/main.go:26-31
        printf("usage: -id 0 -f config.json -s . -m 2 -l -v\n\n");
        return 0;
    }
    std::ifstream in(file);
    if (!in) { fprintf(stderr, "failed to load json file: %s\n", file.c_str()); return 1; }
    std::stringstream ss;
    ss << in.rdbuf();
    int n_dev = dissem_device_count();
    if (n_dev <= 0) { fprintf(stderr, "no CUDA device: this engine has no CPU path\n"); return 1; }
    if (gpus > n_dev) gpus = n_dev;
    int devs[4096];
    for (int i = 0; i < 4096; ++i) devs[i] = i % gpus;
    std::string opts = "{\"batched\":" + std::to_string(batched) + ",\"time_scale\":" + std::to_string(time_scale) +
                       ",\"flags\":3,\"seed\":24301,\"pace\":" + std::to_string(pace) + ",\"dump_tables\":" + std::to_string(dump_tables);
    if (save_disk) {
        std::string esc;
        for (char ch : storage) { if (ch == '"' || ch == '\\') esc += '\\'; esc += ch; }
        opts += ",\"storage\":\"" + esc + "\"";
    }
    opts += "}";
    printf("launching leader...\n[id: %d, filename: %s, storagePath: %s, mode: %d]\n", id, file.c_str(), storage.c_str(), mode);
    char *out = nullptr;
    int rc = dissem_run_json(ss.str().c_str(), mode, devs, opts.c_str(), &out);
    if (rc) { fprintf(stderr, "leader failed: %s\n", dissem_last_error()); return 1; }
    // pull time_to_deliver_ms out of the report
    const char *p = strstr(out, "\"time_to_deliver_ms\":");
    double ms = p ? atof(p + 21) : 0.0;
    printf("Time to deliver: %.6fms\n", ms);                           // cmd/main.go:168
    if (verbose) printf("%s\n", out);
    dissem_free(out);
    return 0;
}
