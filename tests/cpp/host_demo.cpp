// Uses the C++ host mirror (include/sr_renderer.hpp) the way App.swift:153-188 uses Renderer: build a RenderPass, call
// render(renderPass:), read the caller-owned images.  KAT-1 of SURVEY.md §4.  Exit code 0 = all values match.
#include <cmath>
#include <cstdio>
#include <vector>
#include "sr_renderer.hpp"

int main(int argc, char** argv) {
    if (argc > 1 && std::string(argv[1]) == "--link-only") { std::printf("abi %d\n", sr_version()); return 0; }
    const int W = 512, H = 512;
    std::vector<sr::Pixel> color(W * H);
    std::vector<float> depth(W * H);
    sr::RenderPass rp{{color.data(), W, H, W * 4}, {depth.data(), W, H, W * 4}, {}, {}, sr::PrimitiveType::triangle, {}};
    rp.vertices = {{{0, 1, 0, 0}, {1, 0, 0, 0}}, {{1, 0, 0, 0}, {0, 1, 0, 0}}, {{-1, 0, 0, 0}, {0, 0, 1, 0}}};
    rp.indices = {0, 1, 2};
    try {
        sr::CudaRenderer renderer(0);
        if (renderer.render(rp) != 0) return 2;
    } catch (const std::exception& e) { std::fprintf(stderr, "%s\n", e.what()); return 3; }
    long covered = 0;
    for (auto& p : color) covered += p.a != 0;
    auto px = [&](int x, int y) { return color[y * W + x]; };
    bool ok = covered == 65537 && px(256, 0).r == 255 && px(256, 128).b == 63 && px(256, 128).g == 63 && px(256, 128).r == 127 &&
              px(0, 256).b == 255 && px(300, 100).b == 27 && px(300, 100).g == 71 && px(300, 100).r == 155;
    for (float d : depth) ok = ok && std::isinf(d) && d > 0;
    // legacy 2-D API on the same image: a 3-row x 5-column rect (w drives y, Renderer.swift:378) and a 2-D triangle
    try {
        sr::CudaRenderer r2(0);
        r2.drawRect(2, 1, 3, 5, sr::Pixel{9, 8, 7, 6}, rp.colorBuffer);
        r2.drawLine(0, 500, 8, 504, sr::Pixel{5, 6, 7, 8}, rp.colorBuffer);
    } catch (const std::exception& e) { std::fprintf(stderr, "%s\n", e.what()); return 4; }
    ok = ok && px(2, 1).b == 9 && px(6, 3).a == 6 && px(7, 1).a == 0 && px(2, 4).a == 0 && px(0, 500).g == 6 && px(8, 504).a == 0;
    std::printf("covered=%ld ok=%d\n", covered, (int)ok);
    return ok ? 0 : 1;
}
