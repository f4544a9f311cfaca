// args.hpp -- the spotfinder command line (same flags, defaults and error behaviour as
// spotfinder/spotfinder.cc:258-403, src/ffs/arg_parser.cc:35-70 and
// src/ffs/cuda_arg_parser.cc:28-61; the argparse library is not available offline).
#pragma once
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <optional>
#include <string>
#include <vector>

struct SpotfinderArgs {
    std::string file;            // FILE.nxs | /dev/shm stream directory | *.cbf template
    bool sample = false;
    bool verbose = false;
    int device = 0;
    uint32_t threads = 1;
    bool validate = false;
    std::optional<uint32_t> images;
    bool writeout = false;
    uint32_t min_spot_size = 3;
    uint32_t min_spot_size_3d = 3;
    float max_peak_centroid_separation = 2.0f;
    uint32_t start_index = 0;
    float timeout = 30.0f;
    int pipe_fd = -1;
    std::string algorithm = "dispersion";
    float dmin = -1.f, dmax = -1.f;
    std::optional<float> wavelength;
    std::optional<std::string> detector;
    bool save_h5 = false;
    bool output_for_index = false;
    bool list_devices = false;
    std::vector<int> devices;    // --devices 0,1,2 | 0-7: frame-wise sharding over several GPUs (extension)
};

inline void print_usage(const char* prog) {
    printf(
      "Usage: %s [--help] [--version] [-v] [-d DEVICE] [--list-devices] [-n THREADS] [--validate]\n"
      "       [--images N] [--writeout] [--min-spot-size N] [--min-spot-size-3d N]\n"
      "       [--max-peak-centroid-separation F] [--start-index N] [-t SECONDS] [-fd FD]\n"
      "       [-a dispersion|dispersion_extended] [--dmin F] [--dmax F] [-w ANGSTROM]\n"
      "       [--detector JSON] [-h5] [--output-for-index] [--devices LIST] [--sample | FILE.nxs]\n\n"
      "Positional: FILE.nxs  Path to the Nexus file, or a /dev/shm Eiger stream directory\n",
      prog);
}

// Returns false (after printing the error) on a usage error: the caller exits 1.
inline bool parse_args(int argc, char** argv, const char* version, SpotfinderArgs& a) {
    std::vector<std::string> argvv(argv + 1, argv + argc);
    // extra arguments, one per line, from ./common.args if present (arg_parser.cc:58-70)
    {
        std::ifstream f("common.args");
        std::string line;
        while (f && std::getline(f, line))
            if (!line.empty()) argvv.push_back(line);
    }
    if (const char* t = getenv("SPOTFINDER_TIMEOUT")) a.timeout = (float)atof(t);  // spotfinder.cc:293-301
    auto need = [&](size_t& i, const std::string& flag) -> const std::string* {
        if (i + 1 >= argvv.size()) {
            printf("Error: %s: expected 1 argument(s). 0 provided.\n", flag.c_str());
            return nullptr;
        }
        return &argvv[++i];
    };
    try {
        for (size_t i = 0; i < argvv.size(); ++i) {
            const std::string& s = argvv[i];
            const std::string* v = nullptr;
            if (s == "-h" || s == "--help") { print_usage(argv[0]); exit(0); }
            else if (s == "--version") { printf("%s\n", version); exit(0); }
            else if (s == "-v" || s == "--verbose") a.verbose = true;
            else if (s == "--list-devices") a.list_devices = true;
            else if (s == "--sample") a.sample = true;
            else if (s == "--validate") a.validate = true;
            else if (s == "--writeout") a.writeout = true;
            else if (s == "-h5" || s == "--save-h5") a.save_h5 = true;
            else if (s == "--output-for-index") a.output_for_index = true;
            else if (s == "-d" || s == "--device") { if (!(v = need(i, s))) return false; a.device = std::stoi(*v); }
            else if (s == "-n" || s == "--threads") { if (!(v = need(i, s))) return false; a.threads = (uint32_t)std::stoul(*v); }
            else if (s == "--images") { if (!(v = need(i, s))) return false; a.images = (uint32_t)std::stoul(*v); }
            else if (s == "--min-spot-size") { if (!(v = need(i, s))) return false; a.min_spot_size = (uint32_t)std::stoul(*v); }
            else if (s == "--min-spot-size-3d") { if (!(v = need(i, s))) return false; a.min_spot_size_3d = (uint32_t)std::stoul(*v); }
            else if (s == "--max-peak-centroid-separation") { if (!(v = need(i, s))) return false; a.max_peak_centroid_separation = std::stof(*v); }
            else if (s == "--start-index") { if (!(v = need(i, s))) return false; a.start_index = (uint32_t)std::stoul(*v); }
            else if (s == "-t" || s == "--timeout") { if (!(v = need(i, s))) return false; a.timeout = std::stof(*v); }
            else if (s == "-fd" || s == "--pipe_fd") { if (!(v = need(i, s))) return false; a.pipe_fd = std::stoi(*v); }
            else if (s == "-a" || s == "--algorithm") { if (!(v = need(i, s))) return false; a.algorithm = *v; }
            else if (s == "--dmin") { if (!(v = need(i, s))) return false; a.dmin = std::stof(*v); }
            else if (s == "--dmax") { if (!(v = need(i, s))) return false; a.dmax = std::stof(*v); }
            else if (s == "-w" || s == "-\xCE\xBB" || s == "--wavelength") { if (!(v = need(i, s))) return false; a.wavelength = std::stof(*v); }
            else if (s == "--detector") { if (!(v = need(i, s))) return false; a.detector = *v; }
            else if (s == "--devices") {  // "0,1,2" and/or "0-7"
                if (!(v = need(i, s))) return false;
                size_t pos = 0;
                while (pos < v->size()) {
                    size_t end = v->find(',', pos);
                    if (end == std::string::npos) end = v->size();
                    const std::string tok = v->substr(pos, end - pos);
                    const size_t dash = tok.find('-');
                    const int lo = std::stoi(tok.substr(0, dash));
                    const int hi = dash == std::string::npos ? lo : std::stoi(tok.substr(dash + 1));
                    for (int d = lo; d <= hi; ++d) a.devices.push_back(d);
                    pos = end + 1;
                }
            }
            else if (!s.empty() && s[0] == '-' && s.size() > 1 && !std::isdigit((unsigned char)s[1])) {
                printf("Error: Unknown argument: %s\n", s.c_str());
                return false;
            } else if (a.file.empty()) a.file = s;
            else {
                printf("Error: Maximum number of positional arguments exceeded\n");
                return false;
            }
        }
    } catch (const std::exception& e) {
        printf("Error: could not parse argument value (%s)\n", e.what());
        return false;
    }
    if (a.list_devices) return true;
    if (!a.file.empty() && a.sample) {
        printf("Error: Argument '--sample' not allowed with 'FILE.nxs'\n");
        return false;
    }
    if (a.file.empty() && !a.sample && !getenv("H5READ_IMPLICIT_SAMPLE")) {  // spotfinder.cc:266-282
        printf("Error: One of the arguments '--sample' or 'FILE.nxs' is required\n");
        return false;
    }
    return true;
}
