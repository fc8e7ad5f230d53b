// grt_render -- headless stand-in for the reference's viewer main loop (src/main.cpp:9-131): same -p/--ply,
// --width, --height flags and defaults, same object wiring and call order; writes a PPM instead of opening a window.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "GaussianTracer.h"

int main(int argc, char** argv) {
    std::string ply = "../data/train.ply", out = "frame.ppm";
    unsigned width = 1280, height = 720, deg = 0, frames = 1;
    bool fisheye = false, plane = false, sphere = false, device_ingest = false;
    int type = MIRROR;
    for (int i = 1; i < argc; ++i) {
        auto next = [&](const char* flag) -> const char* {
            if (i + 1 >= argc) { fprintf(stderr, "%s needs a value\n", flag); exit(2); }
            return argv[++i];
        };
        if (!strcmp(argv[i], "-p") || !strcmp(argv[i], "--ply")) ply = next("--ply");
        else if (!strcmp(argv[i], "--width")) width = (unsigned)atoi(next("--width"));
        else if (!strcmp(argv[i], "--height")) height = (unsigned)atoi(next("--height"));
        else if (!strcmp(argv[i], "--sh-degree")) deg = (unsigned)atoi(next("--sh-degree"));
        else if (!strcmp(argv[i], "--frames")) frames = (unsigned)atoi(next("--frames"));
        else if (!strcmp(argv[i], "--type")) type = atoi(next("--type"));
        else if (!strcmp(argv[i], "--out")) out = next("--out");
        else if (!strcmp(argv[i], "--fisheye")) fisheye = true;
        else if (!strcmp(argv[i], "--plane")) plane = true;
        else if (!strcmp(argv[i], "--sphere")) sphere = true;
        else if (!strcmp(argv[i], "--device-ingest")) device_ingest = true;
        else if (!strcmp(argv[i], "-h") || !strcmp(argv[i], "--help")) {
            printf("usage: grt_render -p scene.ply [--width W --height H --sh-degree 0..3 --fisheye --plane --sphere --device-ingest --type 0|1|2 --frames N --out frame.ppm|frame.png]\n");
            return 0;
        } else { fprintf(stderr, "unknown flag %s\n", argv[i]); return 2; }
    }
    try {
        GaussianTracer tracer(ply, 0, device_ingest);
        tracer.setSize(width, height);
        tracer.initializeOptix();
        CUDAOutputBuffer output_buffer(tracer.context(), (int32_t)width, (int32_t)height);
        output_buffer.setStream(tracer.stream);
        Camera camera;  // GUI defaults (src/gui.cpp:50-56)
        float3 eye; eye.x = 0.0f; eye.y = 0.0f; eye.z = 3.0f;
        float3 up; up.x = 0.0f; up.y = 1.0f; up.z = 0.0f;
        camera.setEye(eye);
        camera.setLookat(tracer.getGaussianCenter());
        camera.setUp(up);
        camera.setFovY(60.0f);
        bool changed = true;
        tracer.updateCamera(camera, changed);
        tracer.params.sh_degree_max = deg;
        tracer.params.mode_fisheye = fisheye;
        tracer.setRenderType((unsigned)type);
        if (plane) tracer.createPlane();
        if (sphere) tracer.createSphere();
        double total_ms = 0.0;
        for (unsigned f = 0; f < frames; ++f) {
            const auto t0 = std::chrono::steady_clock::now();
            tracer.render(output_buffer);
            total_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        }
        printf("rendered %u frame(s) %ux%u: %.3f ms/frame wall, last kernel %.3f ms\n", frames, width, height, total_ms / frames,
               tracer.lastKernelMs());
        if (out.size() > 4 && out.substr(out.size() - 4) == ".png") output_buffer.writePNG(out.c_str());
        else output_buffer.writePPM(out.c_str());
        printf("wrote %s\n", out.c_str());
    } catch (const std::exception& e) {
        fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
    return 0;
}
