// TEST-ONLY host driver: runs the K6/K7 device code path (contour_core.cuh, which is
// __host__ __device__) sequentially on the CPU so that candidate selection, border following,
// TC89_L1, polygon measurement and the crack-parity externality test can be fuzzed against
// cv2 on a box without a GPU.  Not part of the product library.
#include <algorithm>
#include <cstdlib>
#include <vector>
#include <vector_types.h>
#include <vector_functions.h>

#include "../autobub3hs_b200/csrc/contour_core.cuh"
#include "../include/abub_cuda.h"

using namespace abub_contour;

extern "C" int host_find_contours(const uint8_t* tz, int thr, int rows, int cols, abub_blob* out, int cap) {
    Pix im{tz, thr, rows, cols};
    std::vector<int> cands;
    for (int y = 0; y < rows; y++)
        for (int x = 0; x < cols; x++)
            if (im.fg(x, y) && !im.fg(x - 1, y) && !im.fg(x - 1, y - 1) && !im.fg(x, y - 1) && !im.fg(x + 1, y - 1))
                cands.push_back(y * cols + x);
    std::sort(cands.rbegin(), cands.rend());
    struct Acc { int first_px, len; TraceResult tr; std::vector<uint8_t> mem; int L; PolyMeasure pm; };
    std::vector<Acc> acc;
    for (int p : cands) {
        TraceResult t1 = trace_border(im, p % cols, p / cols, nullptr, 1 << 28);
        if (!t1.accepted) continue;
        Acc a;
        a.first_px = p; a.len = t1.len; a.tr = t1;
        a.mem.resize(ChainView::bytes(t1.len) + 16);
        ChainView cv(a.mem.data(), t1.len);
        if (t1.len == 0) {
            cv.P[0] = make_short2((short)(p % cols), (short)(p / cols));
            cv.removed[0] = 0;
            a.L = 1;
        } else {
            trace_border(im, p % cols, p / cols, &cv, 1 << 28);
            a.L = approx_tc89_l1(cv, t1.len);
        }
        a.pm = measure_alive(cv, a.L);
        acc.push_back(std::move(a));
    }
    int n = 0;
    for (size_t i = 0; i < acc.size(); i++) {
        const int qx = acc[i].first_px % cols, qy = acc[i].first_px / cols;
        bool nested = false;
        for (size_t j = 0; j < acc.size() && !nested; j++) {
            if (j == i || acc[j].len == 0) continue;
            const TraceResult& t = acc[j].tr;
            if (qx <= t.minx || qx >= t.maxx || qy <= t.miny || qy >= t.maxy) continue;
            ChainView cv(acc[j].mem.data(), acc[j].len);
            if (inside_outer_border(cv, acc[j].len, qx, qy)) nested = true;
        }
        if (nested) continue;
        if (n < cap) {
            abub_blob& b = out[n];
            const PolyMeasure& m = acc[i].pm;
            b.k = 0; b.x = m.x; b.y = m.y; b.w = m.w; b.h = m.h; b.npts = m.npts; b.first_px = acc[i].first_px;
            b.reserved = 0; b.sum_x = m.sum_x; b.sum_y = m.sum_y; b.area = m.area; b.m00 = m.m00; b.m10 = m.m10; b.m01 = m.m01;
        }
        n++;
    }
    return n;
}
