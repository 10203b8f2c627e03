// shim/shim_driver.cpp -- TEST INFRASTRUCTURE: the reference's headless main loop (src/main.cpp:106-237) reduced to what
// touches the render path: Scene(file) -> InitDataContainer -> PathtraceInit -> Pathtrace(NULL, 0, 1..n) -> PathtraceFree,
// against the reference's own headers and its own Scene loader, with pathtrace_wavefront:: provided by the shim
// (shim/pathtrace_wavefront_ptb200.cpp) instead of src/pathtrace_wavefront.cu.  Dumps the accumulators so that a test can
// compare them bit for bit with refdump's (the reference's own render of the same file).
//   shimdump SCENE.json OUTDIR ITERS [--pbo]
#include "pathtrace_wavefront.h"
#include "scene.h"
#include "ptb200.h"

#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>
#include <sys/stat.h>

const char* g_scene_file = nullptr;                 // what the shim's "file" mode reads
bool g_enableVisualization = false;                 // extern of the reference's render TU (src/pathtrace_wavefront.cu:41); kept for link parity
namespace pathtrace_wavefront { ptb_renderer* ShimRenderer(); }

static void dump(const std::string& path, const void* p, size_t bytes) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) { fprintf(stderr, "shimdump: cannot write %s\n", path.c_str()); exit(2); }
    fwrite(p, 1, bytes, f);
    fclose(f);
}

int main(int argc, char** argv) {
    if (argc < 4) { fprintf(stderr, "usage: shimdump SCENE.json OUTDIR ITERS [--pbo]\n"); return 1; }
    g_scene_file = argv[1];
    const std::string out = argv[2];
    const int iters = atoi(argv[3]);
    const bool use_pbo = argc > 4 && std::string(argv[4]) == "--pbo";
    mkdir(out.c_str(), 0777);
    cudaSetDevice(0);                               // src/main.cpp:168
    Scene* scene = new Scene(argv[1]);              // the reference's own loader
    GuiDataContainer gui;
    gui.DenoiserOn = false;
    gui.ShowBVH = false;
    pathtrace_wavefront::InitDataContainer(&gui);
    pathtrace_wavefront::PathtraceInit(scene);      // src/main.cpp:33-40 (InitPathTracer)
    const int P = scene->state.camera.resolution.x * scene->state.camera.resolution.y;
    uchar4* pbo = nullptr;
    if (use_pbo) cudaMalloc(&pbo, sizeof(uchar4) * P);
    for (int i = 1; i <= iters; i++) pathtrace_wavefront::Pathtrace(pbo, 0, i);   // src/main.cpp:211-224
    std::vector<float> direct(3 * (size_t)P), indirect(3 * (size_t)P);
    if (ptb_get_accumulators(pathtrace_wavefront::ShimRenderer(), direct.data(), indirect.data()) != PTB_OK) {
        fprintf(stderr, "shimdump: %s\n", ptb_last_error());
        return 3;
    }
    dump(out + "/direct.bin", direct.data(), direct.size() * 4);
    dump(out + "/indirect.bin", indirect.data(), indirect.size() * 4);
    if (pbo) {
        std::vector<unsigned char> px(4 * (size_t)P);
        cudaMemcpy(px.data(), pbo, px.size(), cudaMemcpyDeviceToHost);
        dump(out + "/pbo.bin", px.data(), px.size());
        cudaFree(pbo);
    }
    pathtrace_wavefront::PathtraceFree();
    delete scene;
    printf("SHIMDUMP {\"iters\": %d, \"pixels\": %d, \"mrays_per_s\": %.3f}\n", iters, P, gui.MraysPerSec);
    return 0;
}
