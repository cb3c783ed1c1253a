// pt_render.cpp — headless stand-in for the reference's App/Program (App.cs:15-50, Program.cs:3-7):
// creates the renderer at the window size, renders, and writes the image to a PPM file (the
// "image output" the reference only ever sent to a swapchain).
//   pt_render [--size WxH] [--spp N] [--seed S] [--out file.ppm] [--frames K]
// Build: g++ -O2 -std=c++17 pt_render.cpp -L.. -lptb200 -Wl,-rpath,'$ORIGIN/..' -o pt_render
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "renderer.hpp"

int main(int argc, char **argv)
{
    ptb_host::Window win;
    unsigned spp = 0, seed = 0, frames = 1;
    std::string out = "frame.ppm";
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        if (a == "--size" && i + 1 < argc) {
            if (sscanf(argv[++i], "%ux%u", &win.width, &win.height) != 2) return fprintf(stderr, "bad --size\n"), 2;
        } else if (a == "--spp" && i + 1 < argc) spp = strtoul(argv[++i], nullptr, 0);
        else if (a == "--seed" && i + 1 < argc) seed = strtoul(argv[++i], nullptr, 0);
        else if (a == "--frames" && i + 1 < argc) frames = strtoul(argv[++i], nullptr, 0);
        else if (a == "--out" && i + 1 < argc) out = argv[++i];
        else return fprintf(stderr, "usage: pt_render [--size WxH] [--spp N] [--seed S] [--frames K] [--out file.ppm]\n"), 2;
    }
    try {
        ptb_host::Renderer r(win, 0, seed);
        r.Init();
        const auto t0 = std::chrono::steady_clock::now();
        for (unsigned f = 0; f < frames; ++f) {
            if (spp == 0) {
                r.Render(1.0f / 60.0f);                     // the reference's per-frame path
            } else {                                        // progressive: spp jittered samples, resolve
                ptb_context *c = r.Handle();
                if (ptb_render_resolve(c, f * spp, spp, 1, 0) || ptb_wait(c))      // accumulate + resolve, one launch
                    throw std::runtime_error(ptb_last_error(c));
            }
        }
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        const auto img = r.ReadImage();
        FILE *fp = fopen(out.c_str(), "wb");
        if (!fp) throw std::runtime_error("cannot open " + out);
        fprintf(fp, "P6\n%u %u\n255\n", win.width, win.height);
        for (size_t i = 0; i < (size_t)win.width * win.height; ++i) fwrite(&img[4 * i], 1, 3, fp);
        fclose(fp);
        // the only "metric" the reference prints is a per-frame FPS line (Renderer.cs:1046)
        printf("%ux%u, %u frame(s), %.3f ms/frame -> %s\n", win.width, win.height, frames, ms / frames, out.c_str());
    } catch (const std::exception &e) {
        fprintf(stderr, "pt_render: %s\n", e.what());
        return 1;
    }
    return 0;
}
