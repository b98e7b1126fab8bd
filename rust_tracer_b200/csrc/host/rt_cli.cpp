// rt_cli -- host driver with the reference's command line (src/cli.rs:34-175) and the
// orchestration of src/main.rs:27-291, on top of the C ABI of include/rt_b200.h.  Same flags,
// same stage printouts, same output file naming (./output/<unix-seconds>.png, main.rs:75-79);
// the render itself happens in librt_b200.so on the GPU (no CPU fallback).
//
//   rt_cli [-w W] [-h H] [-d DEPTH] [--method basic|rayforest] [--stats] [-t] [-i]
//          [--scene default|s1k|m10k] [--scene-file FILE] [--dump-scene FILE] [--seed N]
//          [--spp 1|4] [--out FILE] [--device N] [--gpus N]
//          [bench [-n RUNS] [-f]]
// --gpus N renders Basic mode on the first N CUDA devices with ONE rt_multi_render call per frame
// (the library splits the frame into interleaved row tiles and drives the devices itself).
// --scene-file renders a scene description file (include/rt_b200_host.h) instead of a built-in
// scene; --dump-scene writes the chosen scene as such a file and exits (no GPU needed).
// --gui needs a display server (src/gui.rs, GTK): unavailable here, it is rejected.
// The flags after --stats have no counterpart in the reference (its scene is compiled in).
#include <sys/stat.h>

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <string>
#include <vector>

#include "../../../include/rt_b200.h"
#include "../../../include/rt_b200_host.h"

namespace {

// ---- a minimal PNG writer: 8-bit RGB, stored (uncompressed) deflate blocks -------------------
// Stands in for image::ImageBuffer::save (src/bmp.rs:16-18): the pixels are what is checked,
// not the encoder.
uint32_t crc_table[256];
void crc_init() {
    for (uint32_t n = 0; n < 256; ++n) {
        uint32_t c = n;
        for (int k = 0; k < 8; ++k) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
        crc_table[n] = c;
    }
}
uint32_t crc_update(uint32_t c, const uint8_t *p, size_t n) {
    for (size_t i = 0; i < n; ++i) c = crc_table[(c ^ p[i]) & 0xFF] ^ (c >> 8);
    return c;
}
void put32(std::vector<uint8_t> &v, uint32_t x) {
    v.push_back(x >> 24); v.push_back(x >> 16); v.push_back(x >> 8); v.push_back(x);
}
void chunk(std::vector<uint8_t> &out, const char *type, const std::vector<uint8_t> &data) {
    put32(out, (uint32_t)data.size());
    size_t at = out.size();
    out.insert(out.end(), type, type + 4);
    out.insert(out.end(), data.begin(), data.end());
    put32(out, crc_update(0xFFFFFFFFu, out.data() + at, out.size() - at) ^ 0xFFFFFFFFu);
}
bool write_png(const std::string &path, const uint8_t *rgb, int w, int h) {
    crc_init();
    std::vector<uint8_t> raw;
    raw.reserve((size_t)h * (3 * (size_t)w + 1));
    for (int v = 0; v < h; ++v) {
        raw.push_back(0); // filter: none
        raw.insert(raw.end(), rgb + (size_t)v * w * 3, rgb + (size_t)(v + 1) * w * 3);
    }
    std::vector<uint8_t> z = {0x78, 0x01};
    uint32_t a = 1, b = 0;
    for (size_t off = 0; off < raw.size() || off == 0; off += 65535) {
        size_t n = std::min<size_t>(65535, raw.size() - off);
        z.push_back(off + n >= raw.size() ? 1 : 0);
        z.push_back(n & 0xFF); z.push_back(n >> 8); z.push_back(~n & 0xFF); z.push_back((~n >> 8) & 0xFF);
        z.insert(z.end(), raw.begin() + off, raw.begin() + off + n);
        for (size_t i = 0; i < n; ++i) { a = (a + raw[off + i]) % 65521; b = (b + a) % 65521; }
        if (raw.empty()) break;
    }
    put32(z, (b << 16) | a);
    std::vector<uint8_t> out = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A}, ihdr;
    put32(ihdr, (uint32_t)w); put32(ihdr, (uint32_t)h);
    ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
    chunk(out, "IHDR", ihdr);
    chunk(out, "IDAT", z);
    chunk(out, "IEND", {});
    FILE *f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    bool ok = std::fwrite(out.data(), 1, out.size(), f) == out.size();
    return std::fclose(f) == 0 && ok;
}

// ---- Config (src/cli.rs:3-32) ------------------------------------------------------------------
struct Config {
    long width = 512, height = 512, depth = 8; // defaults: cli.rs:43,51,59
    bool to_terminal = false, gui = false, interactive = false, print_forest_stats = false;
    bool rayforest = false;                    // Method::{Basic, RayForest}, default basic (cli.rs:66)
    bool bench = false;
    int runs = 10;                             // cli.rs:93
    bool filter_mode = false;
    // extensions
    std::string scene = "default", scene_file, dump_scene, out;
    uint64_t seed = 0;
    int spp = 1, device = 0, gpus = 1;
};

[[noreturn]] void die(const std::string &msg) { // the reference panics (cli.rs:123 ff.)
    std::fprintf(stderr, "rt_cli: %s\n", msg.c_str());
    std::exit(101);
}

long parse_int(const char *s, const char *what) {
    char *end = nullptr;
    long v = std::strtol(s, &end, 10);
    if (!s[0] || *end || v < 0) die(std::string("Expected integer for ") + what);
    return v;
}

Config parse_args(int argc, char **argv) {
    Config c;
    bool in_bench = false;
    auto value = [&](int &i) -> const char * {
        if (i + 1 >= argc) die(std::string("missing value for ") + argv[i]);
        return argv[++i];
    };
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        if (in_bench) {
            if (a == "-n" || a == "--runs") c.runs = (int)parse_int(value(i), "number of runs");
            else if (a == "-f" || a == "--filter") c.filter_mode = true;
            else die("unexpected argument to bench: " + a);
        } else if (a == "bench") { c.bench = in_bench = true; }
        else if (a == "-w" || a == "--width") c.width = parse_int(value(i), "width");
        else if (a == "-h" || a == "--height") c.height = parse_int(value(i), "height"); // -h is height (cli.rs:47-52)
        else if (a == "-d" || a == "--depth") c.depth = parse_int(value(i), "depth");
        else if (a == "--method") {
            std::string m = value(i);
            for (auto &ch : m) ch = (char)std::tolower((unsigned char)ch); // cli.rs:136
            if (m == "rayforest") c.rayforest = true;
            else if (m == "basic") c.rayforest = false;
            else die("Unexpected value provided for `--method`: " + m);
        } else if (a == "-t" || a == "--to-terminal") c.to_terminal = true;
        else if (a == "-i" || a == "--interactive") c.interactive = true;
        else if (a == "--stats") c.print_forest_stats = true;
        else if (a == "-g" || a == "--gui") c.gui = true;
        else if (a == "--scene") c.scene = value(i);
        else if (a == "--scene-file") c.scene_file = value(i);
        else if (a == "--dump-scene") c.dump_scene = value(i);
        else if (a == "--seed") c.seed = (uint64_t)parse_int(value(i), "seed");
        else if (a == "--spp") c.spp = (int)parse_int(value(i), "spp");
        else if (a == "--out") c.out = value(i);
        else if (a == "--device") c.device = (int)parse_int(value(i), "device");
        else if (a == "--gpus") c.gpus = (int)parse_int(value(i), "gpus");
        else die("unexpected argument: " + a);
    }
    return c;
}

#define CHECK(call)                                                                  \
    do {                                                                             \
        if ((call) != RT_OK) die(std::string(#call ": ") + rt_last_error());          \
    } while (0)

using Clock = std::chrono::steady_clock;
long long ms_since(Clock::time_point t0) {
    return std::chrono::duration_cast<std::chrono::milliseconds>(Clock::now() - t0).count();
}
long long ns_since(Clock::time_point t0) {
    return std::chrono::duration_cast<std::chrono::nanoseconds>(Clock::now() - t0).count();
}

void enter_to_proceed() { // main.rs:125-131
    std::printf("Enter To Proceed: ");
    std::fflush(stdout);
    int ch;
    while ((ch = std::getchar()) != '\n' && ch != EOF) {}
}

rt_render_params params(const Config &c, long depth) {
    rt_render_params p;
    std::memset(&p, 0, sizeof(p));
    p.depth = (int32_t)depth; p.spp = c.spp; p.seed = c.seed;
    p.rank = 0; p.world = 1; p.tile_rows = 8; p.flags = 0;
    return p;
}

// terminal::draw (render.rs:188-219) of draw_to_terminal's 100x50 depth-5 re-render (render.rs:21-29)
void draw_to_terminal(const Config &c, rt_scene *scene) {
    const int w = 100, h = 50;
    rt_camera cam;
    rt_camera_default(w, h, &cam);
    rt_render_params p = params(c, 5);
    std::vector<float> f((size_t)w * h * 3);
    std::vector<uint8_t> u((size_t)w * h * 3);
    CHECK(rt_render(scene, &cam, &p, f.data(), u.data(), nullptr));
    for (int v = 0; v < h; ++v) {
        for (int x = 0; x < w; ++x) {
            size_t i = ((size_t)v * w + x) * 3;
            if (f[i] == 0.f && f[i + 1] == 0.f && f[i + 2] == 0.f) std::printf("\x1b[37m.");
            else std::printf("\x1b[38;2;%d;%d;%dmX", u[i], u[i + 1], u[i + 2]);
        }
        std::printf("\n");
    }
    std::printf("\x1b[0m");
}

// render_scene_basic (main.rs:244-261); multi != null: on all its devices with one call
std::vector<uint8_t> render_scene_basic(const Config &c, rt_scene *scene, rt_multi *multi) {
    rt_camera cam;
    rt_camera_default((int32_t)c.width, (int32_t)c.height, &cam);
    rt_render_params p = params(c, c.depth);
    std::vector<uint8_t> u8((size_t)c.width * c.height * 3);
    auto t0 = Clock::now();
    if (multi) CHECK(rt_multi_render(multi, &cam, &p, nullptr, u8.data(), nullptr));
    else CHECK(rt_render(scene, &cam, &p, nullptr, u8.data(), nullptr));
    std::printf("render_scene: %lldms\n", ms_since(t0));
    if (c.to_terminal && scene) draw_to_terminal(c, scene);
    return u8;
}

rt_forest *generate_forest(const Config &c, rt_scene *scene) { // main.rs:263-274
    rt_camera cam;
    rt_camera_default((int32_t)c.width, (int32_t)c.height, &cam);
    rt_render_params p = params(c, c.depth);
    p.spp = 1;
    rt_forest *f = nullptr;
    auto t0 = Clock::now();
    CHECK(rt_forest_build(scene, &cam, &p, &f, nullptr));
    std::printf("generate_forest: %lldms\n", ms_since(t0));
    return f;
}

std::vector<uint8_t> render_forest(const Config &c, rt_forest *f, const float *ambient) { // main.rs:276-291
    std::vector<uint8_t> u8((size_t)c.width * c.height * 3);
    auto t0 = Clock::now();
    CHECK(rt_forest_eval(f, ambient, nullptr, u8.data(), nullptr));
    std::printf("render_forest: %lldms\n", ms_since(t0));
    return u8;
}

std::string output_path(const Config &c) { // "./output/<unix-seconds>.png" (main.rs:75-79)
    if (!c.out.empty()) return c.out;
    mkdir("./output", 0777);
    return "./output/" + std::to_string((long long)std::time(nullptr)) + ".png";
}

void save(const Config &c, const std::vector<uint8_t> &u8) {
    std::string path = output_path(c);
    if (!write_png(path, u8.data(), (int)c.width, (int)c.height)) die("Failed to save image to disk: " + path);
}

void print_totals(long long ns, int runs) { // main.rs:141-151
    long long ms = ns / 1000000;
    std::printf("Total Time: %lldms | %lldns\n", ms, ns);
    std::printf("Avg Per Op: %gms | %gns\n", (double)((float)ms / (float)runs), (double)((float)ns / (float)runs));
}

} // namespace

int main(int argc, char **argv) {
    Config c = parse_args(argc, argv);
    std::printf("Rendering configuration: Config { width: %ld, height: %ld, depth: %ld, to_terminal: %s, gui: %s, method: %s, "
                "interactive: %s, subcommand: %s, print_forest_stats: %s }\n",
                c.width, c.height, c.depth, c.to_terminal ? "true" : "false", c.gui ? "true" : "false",
                c.rayforest ? "RayForest" : "Basic", c.interactive ? "true" : "false", c.bench ? "Benchmark" : "Normal",
                c.print_forest_stats ? "true" : "false");
    if (c.gui) die("--gui needs a display server (GTK, src/gui.rs): unavailable here; use bench --filter for the re-shade path");
    if (c.width <= 0 || c.height <= 0) die("width and height must be positive");

    std::printf("Create Scene\n");
    rt_scene_desc *desc = nullptr;
    if (!c.scene_file.empty()) CHECK(rt_scene_desc_load(c.scene_file.c_str(), &desc));
    else if (rt_builtin_scene(c.scene.c_str(), 0, &desc) != RT_OK) die("unknown scene: " + c.scene);
    if (!c.dump_scene.empty()) { // host only: no device is touched
        CHECK(rt_scene_desc_save(desc, c.dump_scene.c_str()));
        std::printf("Wrote scene description: %s (%d shapes, %d materials, %d lights)\n", c.dump_scene.c_str(), desc->n_shapes,
                    desc->n_materials, desc->n_lights);
        rt_scene_desc_free(desc);
        return 0;
    }
    rt_scene *scene = nullptr;
    rt_multi *multi = nullptr;
    if (c.gpus > 1) {
        if (c.rayforest) die("--gpus: the ray forest lives on one device; use --method basic");
        int n = rt_device_count();
        if (n < 0) die(std::string("rt_device_count: ") + rt_last_error());
        if (c.gpus > n) die("--gpus " + std::to_string(c.gpus) + ": only " + std::to_string(n) + " CUDA devices");
        std::vector<int32_t> devices;
        for (int d = 0; d < c.gpus; ++d) devices.push_back(d);
        CHECK(rt_multi_create(desc, devices.data(), c.gpus, &multi));
        if (c.to_terminal) CHECK(rt_scene_create(desc, 0, &scene)); // the 100x50 terminal preview: one device
    } else {
        CHECK(rt_scene_create(desc, c.device, &scene));
    }
    std::printf("Done Creating Scene\n");
    const float *ambient = desc->ambient;

    if (!c.bench) { // handle_normal_mode (main.rs:45-123)
        if (c.interactive) enter_to_proceed();
        if (!c.rayforest) {
            std::printf("Rendering in Basic Mode\n");
            auto t0 = Clock::now();
            save(c, render_scene_basic(c, scene, multi));
            std::printf("render_basic_to_file: %lldms\n", ms_since(t0));
        } else {
            std::printf("Rendering in RayForest Mode\nGenerate Forest\n");
            rt_forest *forest = generate_forest(c, scene);
            std::printf("Done Generating Forest\n");
            if (c.print_forest_stats) { // main.rs:88-103
                rt_forest_stats st;
                CHECK(rt_forest_get_stats(forest, &st));
                std::printf("Number of Trees: %llu\nMin Tree Size: %llu\nMax Tree Size: %llu\nMedian Size: %llu\n"
                            "p90 Size: %llu\np95 Size: %llu\np99 Size: %llu\nNumber of Intersections: %llu\n",
                            (unsigned long long)st.num_trees, (unsigned long long)st.smallest_tree,
                            (unsigned long long)st.largest_tree, (unsigned long long)st.median, (unsigned long long)st.p90,
                            (unsigned long long)st.p95, (unsigned long long)st.p99, (unsigned long long)st.num_intersections);
                std::printf("Number of Shapes: %d\nNumber of Intersection Tests: %llu\n", desc->n_shapes,
                            (unsigned long long)desc->n_shapes * (unsigned long long)st.num_intersections);
            }
            if (c.interactive) enter_to_proceed();
            auto t0 = Clock::now();
            save(c, render_forest(c, forest, ambient));
            std::printf("render_forest_to_file: %lldms\n", ms_since(t0));
            rt_forest_destroy(forest);
        }
    } else if (!c.rayforest) { // handle_benchmark_mode, Basic (main.rs:136-151)
        auto t0 = Clock::now();
        for (int i = 0; i < c.runs; ++i) render_scene_basic(c, scene, multi);
        print_totals(ns_since(t0), c.runs);
    } else { // RayForest (main.rs:153-217)
        std::printf("This will benchmark evaluating the complete forest\nRendering in RayForest Mode\nGenerate Forest\n");
        rt_forest *forest = generate_forest(c, scene);
        std::printf("Done Generating Forest\n");
        long long ns;
        if (!c.filter_mode) {
            std::printf("Render full forest\n");
            auto t0 = Clock::now();
            for (int i = 0; i < c.runs; ++i) render_forest(c, forest, ambient);
            ns = ns_since(t0);
        } else {
            std::printf("Render partial forest\n");
            const int32_t shape_id = 1; // find_shape("blue"): the second shape of my_scene.rs (my_scene.rs:62-73)
            std::vector<uint8_t> u8((size_t)c.width * c.height * 3, 0);
            auto t0 = Clock::now();
            for (int i = 0; i < c.runs; ++i)
                CHECK(rt_forest_eval_filtered(forest, ambient, &shape_id, 1, nullptr, u8.data(), nullptr, nullptr));
            ns = ns_since(t0);
            uint64_t tree_count = 0, trees_with = 0;
            CHECK(rt_forest_size(forest, &tree_count));
            CHECK(rt_forest_trees_with(forest, shape_id, &trees_with));
            std::printf("Forest Size: %llu\nTrees Evaluated: %llu\n", (unsigned long long)tree_count,
                        (unsigned long long)trees_with);
            std::printf("%% evaluated: %g\n", (double)(100.f * (float)trees_with / (float)tree_count));
        }
        print_totals(ns, c.runs);
        rt_forest_destroy(forest);
    }
    rt_multi_destroy(multi);
    rt_scene_destroy(scene);
    rt_scene_desc_free(desc);
    return 0;
}
