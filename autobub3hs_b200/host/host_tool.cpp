// host_tool -- test driver for the host-side pieces that need no GPU (tests/test_host_cpu.py):
//   host_tool decode <file> <out.raw>            decode BMP/PNG -> "rows cols\n" on stdout + raw gray bytes
//   host_tool raw-list <run_dir/> <ImageFolder> <event> <cam>     RawParser::ParseAndSortFramesInFolder
//   host_tool raw-events <run_dir/>                               RawParser::GetEventDirLists
//   host_tool zip-list <run.zip> <event> <cam> | zip-events <run.zip> | zip-get <run.zip> <event> <frame> <out.raw>
//   host_tool writer-selftest <out_dir/>         fixed bubble lists through OutputWriter (header + rows)
#include "RankSync.hpp"
#include <time.h>
#include <chrono>
#include <cstring>
#include <cstdint>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <iostream>
#include <regex>

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
        if (cmd == "rank-sync" && argc >= 7) {
            /* rank-sync <base> <rank> <world> <start time or 0> <phase>=<status> ... : the status rendezvous of abub3hs
             * --world N, one phase after the other; prints the agreed value of every phase */
            RankSync rs;
            rs.base = argv[2]; rs.rank = atoi(argv[3]); rs.world = atoi(argv[4]);
            rs.t_start = atoll(argv[5]) ? atoll(argv[5]) : (long long)time(nullptr);
            for (int i = 6; i < argc; i++) {
                const std::string a = argv[i];
                const size_t eq = a.find('=');
                if (eq == std::string::npos) return 1;
                const int got = rs.agree(a.substr(0, eq).c_str(), atoi(a.substr(eq + 1).c_str()));
                printf("%s %d\n", a.substr(0, eq).c_str(), got);
                fflush(stdout);
                if (got < 0) break;
            }
            usleep(300000);
            rs.forget(0);
            return 0;
        }
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
            std::string fast;
            w.appendCameraOutput(fast);                    // batch formatter (to_chars) must print the same bytes
            if (fast != w.formatCameraOutput()) { printf("BATCH FORMATTER DIFFERS\n%s", fast.c_str()); return 1; }
            w.writeCameraOutput();
            for (int c = 0; c < 4; c++) w.stageCameraOutputError(c, -7, 4);
            fast.clear();
            w.appendCameraOutput(fast);
            if (fast != w.formatCameraOutput()) { printf("BATCH FORMATTER DIFFERS\n%s", fast.c_str()); return 1; }
            w.writeCameraOutput();
            printf("%s\n", w.fileName().c_str());
            return 0;
        }
        if (cmd == "writer-bench" && argc == 4) {
            // events/s of the per-event path (format with iostream, open/append/close per event: PICOFormatWriterV4.cpp:287-302)
            // and of the batch path (to_chars into one buffer, one append) on an event with 2 tracked bubbles per camera
            const int nev = atoi(argv[3]);
            std::vector<bubble*> bl;
            for (int c = 0; c < 8; c++) {
                bubble* b = new bubble(frame(100 + c, 50, 4, 3, 7.5, 101.256f + c, 51.5f));
                for (int k = 1; k <= 10; k++) {
                    b->lockThisIteration = false;
                    *b << frame(100 + c, 50 - k, 4 + k, 3 + k, 7.5 + 2.1 * k, 101.3f + c, 51.5f - 1.25f * k);
                }
                bl.push_back(b);
            }
            auto stage = [&](OutputWriter& w, int ev) {
                for (int c = 0; c < 4; c++) w.stageCameraOutput(std::vector<bubble*>{bl[2 * c], bl[2 * c + 1]}, c, 17, ev);
            };
            auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
            const std::string dir = argv[2];
            double t0 = now();
            {
                OutputWriter w(dir, "per_event", 30, 4);
                w.writeHeader();
                for (int ev = 0; ev < nev; ev++) { stage(w, ev); w.writeCameraOutput(); }
            }
            const double t_event = now() - t0;
            t0 = now();
            {
                OutputWriter w(dir, "batch", 30, 4);
                w.writeHeader();
                std::string rows;
                for (int ev = 0; ev < nev; ev++) {
                    stage(w, ev);
                    w.appendCameraOutput(rows);
                    if ((ev + 1) % 1024 == 0) { OutputWriter::appendToFile(w.fileName(), rows); rows.clear(); }
                }
                OutputWriter::appendToFile(w.fileName(), rows);
            }
            const double t_batch = now() - t0;
            printf("{\"events\": %d, \"per_event_events_per_s\": %.0f, \"batch_events_per_s\": %.0f}\n", nev, nev / t_event, nev / t_batch);
            return 0;
        }
        if (cmd == "frame-pattern-fuzz" && argc == 3) {
            // RawParser::FramePathMatches against the std::regex it replaces, on random strings over an alphabet that
            // makes near-misses common (prefixes of "cam<N>", the two suffixes, line breaks)
            const long n = atol(argv[2]);
            uint64_t x = 0x2545F4914F6CDD1Dull;
            auto rnd = [&x]() { x ^= x << 13; x ^= x >> 7; x ^= x << 17; return (uint32_t)(x >> 20); };
            const char alpha[] = "cam0123pngb.m/_\n\r x";
            long bad = 0, hits = 0;
            for (int cam = 0; cam < 4; cam++) {
                const std::regex re("^.*cam" + std::to_string(cam) + ".*(png|bmp)");
                for (long i = 0; i < n; i++) {
                    std::string s;
                    const int len = (int)(rnd() % 14);
                    for (int k = 0; k < len; k++) s += alpha[rnd() % (sizeof(alpha) - 1)];
                    if (rnd() % 4 == 0) s = "cam" + std::to_string(rnd() % 4) + s;
                    if (rnd() % 3 == 0) s += (rnd() % 2) ? "png" : "bmp";
                    if (rnd() % 5 == 0) s = "/data/run/7/Images/" + s;
                    const bool a = std::regex_match(s, re), b = RawParser::FramePathMatches(s, cam);
                    hits += a;
                    if (a != b && bad++ < 5) printf("'%s' cam %d: regex %d matcher %d\n", s.c_str(), cam, (int)a, (int)b);
                }
            }
            printf("%ld strings, %ld matches, %ld mismatches\n", 4 * n, hits, bad);
            return bad ? 1 : 0;
        }
        if (cmd == "format-fuzz" && argc == 3) {
            // the two value formatters on random floats (as double, the promotion `stream << float` does), on halves of the
            // last printed digit, on huge/tiny/non-finite values
            const long n = atol(argv[2]);
            uint64_t x = 0x9E3779B97F4A7C15ull;
            long bad = 0;
            auto check = [&](double v) {
                if (OutputWriter::formatValueFast(v) != OutputWriter::formatValueStream(v)) {
                    if (bad++ < 5) printf("%a: fast '%s' stream '%s'\n", v, OutputWriter::formatValueFast(v).c_str(), OutputWriter::formatValueStream(v).c_str());
                }
            };
            for (long i = 0; i < n; i++) {
                x ^= x << 13; x ^= x >> 7; x ^= x << 17;
                uint32_t bits = (uint32_t)(x >> 17);
                float f;
                memcpy(&f, &bits, 4);
                check((double)f);                                           // any float, incl. NaN/inf/denormals
                check((double)(float)((int)(x % 400000) - 200000) / 8.0);  // multiples of 0.125 -> ties at the 3rd decimal
                check((double)(float)(((int)(x % 4000000) - 2000000) * 0.005f));
                check(((double)(int64_t)(x % 2000001) - 1000000) / 1000.0);
            }
            volatile double z = 0.0;
            const double special[] = {0.0, -0.0, 0.005, 0.015, 0.025, -0.005, 1e300, -1e300, 1e-300, 3.4028234663852886e38,
                                      1.0 / z, -1.0 / z, z / z, -(z / z), 0.125, 0.375, 2.675, 1.005};
            for (double v : special) check(v);
            printf("%ld values, %ld mismatches\n", 4 * n + (long)(sizeof(special) / sizeof(special[0])), bad);
            return bad ? 1 : 0;
        }
    } catch (int code) {
        printf("THROW %d\n", code);
        return 0;
    }
    fprintf(stderr, "bad arguments\n");
    return 1;
}
