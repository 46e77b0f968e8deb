// host_tool -- test driver for the host-side pieces that need no GPU (tests/test_host_cpu.py):
//   host_tool decode <file> <out.raw>            decode BMP/PNG -> "rows cols\n" on stdout + raw gray bytes
//   host_tool raw-list <run_dir/> <ImageFolder> <event> <cam>     RawParser::ParseAndSortFramesInFolder
//   host_tool raw-events <run_dir/>                               RawParser::GetEventDirLists
//   host_tool zip-list <run.zip> <event> <cam> | zip-events <run.zip> | zip-get <run.zip> <event> <frame> <out.raw>
//   host_tool writer-selftest <out_dir/>         fixed bubble lists through OutputWriter (header + rows)
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <iostream>

#include "ImageCodec.hpp"
#include "PICOFormatWriterV4.hpp"
#include "RawParser.hpp"
#include "ZipParser.hpp"

static int dump(const Image8& im, const char* out) {
    if (im.empty()) { printf("FAIL\n"); return 2; }
    printf("%d %d\n", im.rows, im.cols);
    FILE* f = fopen(out, "wb");
    if (!f) return 3;
    fwrite(im.data.data(), 1, im.data.size(), f);
    fclose(f);
    return 0;
}

static BubbleImageFrame frame(int x, int y, int w, int h, double area, float cx, float cy) {
    BubbleImageFrame f;
    f.newPosition = RectI{x, y, w, h};
    f.ContArea = area;
    f.ContRadius = sqrt(area / 3.14159);
    f.MassCentres.x = cx; f.MassCentres.y = cy;
    return f;
}

int main(int argc, char** argv) {
    if (argc < 2) return 1;
    const std::string cmd = argv[1];
    try {
        if (cmd == "decode" && argc == 4) {
            std::vector<uint8_t> b;
            Image8 im;
            if (!abubcodec::read_file(argv[2], b)) { printf("FAIL\n"); return 2; }
            abubcodec::decode(b.data(), b.size(), im);
            return dump(im, argv[3]);
        }
        if (cmd == "raw-list" && argc == 6) {
            RawParser p(argv[2], argv[3], "cam%d_image%u.png");
            std::vector<std::string> c;
            p.ParseAndSortFramesInFolder(argv[4], atoi(argv[5]), c);
            for (auto& s : c) printf("%s\n", s.c_str());
            return 0;
        }
        if (cmd == "raw-events" && argc == 3) {
            RawParser p(argv[2], "/Images/", "cam%d_image%u.png");
            std::vector<std::string> c;
            p.GetEventDirLists(c);
            for (auto& s : c) printf("%s\n", s.c_str());
            return 0;
        }
        if (cmd == "zip-list" && argc == 5) {
            ZipParser p(argv[2], "/Images/", "cam%d_image%u.png");
            ZipParser* q = p.clone();                      // listing through a clone, as the analyzers do
            std::vector<std::string> c;
            q->ParseAndSortFramesInFolder(argv[3], atoi(argv[4]), c);
            for (auto& s : c) printf("%s\n", s.c_str());
            delete q;
            return 0;
        }
        if (cmd == "zip-events" && argc == 3) {
            ZipParser p(argv[2], "/Images/", "cam%d_image%u.png");
            std::vector<std::string> c;
            p.GetEventDirLists(c);
            for (auto& s : c) printf("%s\n", s.c_str());
            return 0;
        }
        if (cmd == "zip-get" && argc == 6) {
            ZipParser p(argv[2], "/Images/", "cam%d_image%u.png");
            Image8 im;
            const int rc = p.GetImage(argv[3], argv[4], im);
            if (rc == -1) { printf("MISSING\n"); return 0; }
            return dump(im, argv[5]);
        }
        if (cmd == "writer-selftest" && argc == 3) {
            OutputWriter w(argv[2], "20210101_0", 30, 4);
            w.writeHeader();
            // cam 0: two bubbles -- one untracked (dZdt = 0/0), one tracked for 12 frames (more than the 10 columns)
            bubble b0(frame(100, 50, 4, 3, 7.5, 101.256f, 51.5f));
            bubble b1(frame(10, 20, 6, 6, 20.0, 12.994999f, 23.005f));
            for (int k = 1; k <= 12; k++) {
                b1.lockThisIteration = false;
                b1 << frame(10 - k, 20 + k, 6 + k, 6 + 2 * k, 20.0 + 3.3 * k, 12.25f - k, 23.125f + 1.5f * k);
            }
            std::vector<bubble*> l0{&b0, &b1};
            w.stageCameraOutput(l0, 0, 17, 3);
            // cam 1: tracked for 3 frames, NaN centroid in one of them
            bubble b2(frame(500, 400, 2, 9, 0.0, 500.5f, 404.0f));
            for (int k = 1; k <= 3; k++) {
                b2.lockThisIteration = false;
                volatile float z = 0.f;
                b2 << frame(500, 400 + k, 3, 9, 0.5 * k, k == 2 ? z / z : 500.125f, 404.0f + k);
            }
            std::vector<bubble*> l1{&b2};
            w.stageCameraOutput(l1, 1, 33, 3);
            w.stageCameraOutputError(2, -3, 3);            // cam 2: no trigger
            std::vector<bubble*> none;
            w.stageCameraOutput(none, 3, 5, 3);            // cam 3: empty list stages -1
            w.writeCameraOutput();
            for (int c = 0; c < 4; c++) w.stageCameraOutputError(c, -7, 4);
            w.writeCameraOutput();
            printf("%s\n", w.fileName().c_str());
            return 0;
        }
    } catch (int code) {
        printf("THROW %d\n", code);
        return 0;
    }
    fprintf(stderr, "bad arguments\n");
    return 1;
}
