#include "L3Localizer.hpp"

#include <math.h>

#include <iostream>

#include "AbubDevice.hpp"
#include "CommonParameters.h"
#include "ImageCodec.hpp"

L3Localizer::L3Localizer(std::string EventID, std::string ImageDir, int CameraNumber, bool nonStopPref,
                         Trainer** TrainedData, std::string MaskDir, Parser* parser)
    : AnalyzerUnit(EventID, ImageDir, CameraNumber, TrainedData, MaskDir, parser) {
    nonStopMode = nonStopPref;
    Level1SuspicionFlag = false;
    numBubbleMultiplicity = 0;
}

L3Localizer::~L3Localizer() {}

void L3Localizer::printBubbleList(void) {
    for (bubble* b : BubbleList) b->printAllXY();
}

// L3Localizer.cpp:971-1012
bool L3Localizer::isInMask(RectI* r, bool bellows) {
    const int xpix = (int)(r->x + r->width / 2.);
    const int ypix = r->y + r->height / 2;
    if (MaskDir == "") return !bellows;                                   // silently skip the mask check
    /* the analyzer keeps its masks once read (cam_mask / bellows_mask members of the reference, :983-984) */
    const Image8*& cached = bellows ? bellows_mask_ : cam_mask_;
    bool& looked = bellows ? bellows_mask_looked_ : cam_mask_looked_;
    if (!looked) { cached = AbubDevice::Get().Mask(CameraNumber, bellows, MaskDir); looked = true; }
    const Image8* mp = cached;
    if (!mp) {
        std::cout << "Mask image not loadable for event " << EventID << " camera " << CameraNumber
                  << "; skipping mask check" << std::endl;
        return !bellows;
    }
    const Image8& mask = *mp;
    if (ypix < 0 || ypix >= mask.rows || xpix < 0 || xpix >= mask.cols) return false;   // cv::Mat::at would read out of bounds
    if ((int)mask.at(ypix, xpix) > 0) return true;
    if (!nonStopMode && !bellows)
        std::cout << "\nCaught a bubble at: (" << xpix << "," << ypix << ") cam " << CameraNumber << ", event " << EventID
                  << ") with pixel value: " << (int)mask.at(ypix, xpix) << " (out of image mask bounds)\n";
    return false;
}

// L3Localizer.cpp:401-419 (genesis: centroid falls back to the mean of the polygon vertices when m00 <= 0)
// and :818-823 (post-trigger: no guard, 0/0 gives NaN which never matches a bubble)
BubbleImageFrame L3Localizer::FrameFromBlob(const abub_blob& b, bool genesis) {
    BubbleImageFrame f;
    f.newPosition.x = b.x; f.newPosition.y = b.y; f.newPosition.width = b.w; f.newPosition.height = b.h;
    f.ContArea = b.area;
    f.ContRadius = sqrt(b.area / 3.14159);
    f.moments.m00 = b.m00; f.moments.m10 = b.m10; f.moments.m01 = b.m01;
    if (!genesis || b.m00 > 0) {
        volatile double m00 = b.m00;
        f.MassCentres.x = (float)(b.m10 / m00);
        f.MassCentres.y = (float)(b.m01 / m00);
    } else {
        f.MassCentres.x = (float)((double)b.sum_x / (double)b.npts);
        f.MassCentres.y = (float)((double)b.sum_y / (double)b.npts);
    }
    return f;
}

void L3Localizer::GenesisFromBlobs(const abub_blob* blobs, int n) {
    /* contours whose box centre lies in the bellows mask are dropped; `largest` counts the others (:273-290) */
    std::vector<const abub_blob*> kept;
    int largest = 0;
    bool all_in_bellows = n > 0;
    for (int i = 0; i < n; i++) {
        RectI r{blobs[i].x, blobs[i].y, blobs[i].w, blobs[i].h};
        if (!isInMask(&r, true)) {
            all_in_bellows = false;
            if (largest < r.width * r.height) largest = r.width * r.height;
            kept.push_back(&blobs[i]);
        }
    }
    std::vector<abub_blob> again;
    if (all_in_bellows) {
        /* :292-390  subtract the bellows movement and see if contours remain.  The contours are found anew (:374) and
         * this time none is dropped for lying in the bellows mask; `largest` runs over all of them (:383-389).  Without
         * a loadable template the binary image is unchanged, i.e. the contours are the ones already found. */
        kept.clear();
        if (BellowsVeto(again)) {
            for (const abub_blob& b : again) kept.push_back(&b);
        } else {
            for (int i = 0; i < n; i++) kept.push_back(&blobs[i]);
        }
        largest = 0;
        for (const abub_blob* b : kept)
            if (largest < b->w * b->h) largest = b->w * b->h;
    }
    for (const abub_blob* b : kept) {                                      // :392-444
        const int BoxArea = b->w * b->h;
        if (BoxArea > 10 || BoxArea >= largest) {
            BubbleImageFrame f = FrameFromBlob(*b, true);
            if (!isInMask(&f.newPosition)) continue;                       // :432
            BubbleList.push_back(new bubble(f));
        }
    }
}

bool L3Localizer::BellowsVeto(std::vector<abub_blob>& again) {
    AbubDevice& dev = AbubDevice::Get();
    std::lock_guard<std::recursive_mutex> g(dev.lock());
    if (!dev.EnsureBellowsTemplate(CameraNumber, MaskDir)) {               // :297-301
        std::cout << "Template image not loadable for event " << EventID << " camera " << CameraNumber
                  << "; cannot veto bellows movement triggers" << std::endl;
        std::cout << "Template path: " << MaskDir + "/cam" + std::to_string(CameraNumber) + "_bellows_template.png" << std::endl;
        return false;
    }
    float pos_trig[2], pos_pre[2];
    /* TrackAFeature on the trigger and the pre-trigger frame (:313-314); a peak on the edge of the correlation map is
     * an out-of-bounds read in the reference -- here it ends the camera like any other exception (-6) */
    dev.Check(abub_bellows_track(dev.ctx(), Slot, CameraNumber, MatTrigFrame, &pos_trig[0], &pos_trig[1], nullptr), "abub_bellows_track");
    dev.Check(abub_bellows_track(dev.ctx(), Slot, CameraNumber, preTrigNumber, &pos_pre[0], &pos_pre[1], nullptr), "abub_bellows_track");
    again.resize(1024);
    int n = 0, status = 0;
    for (;;) {
        const int rc = abub_localize_bellows(dev.ctx(), Slot, CameraNumber, MatTrigFrame, preTrigNumber, loc_thres, pos_trig, pos_pre,
                                             again.data(), (int)again.size(), &n, &status, nullptr);
        if (rc == ABUB_E_CAPACITY && again.size() < (1u << 22)) { again.resize(again.size() * 4); continue; }
        dev.Check(rc, "abub_localize_bellows");
        break;
    }
    if (status != 0)   /* cv::Mat::operator()(Rect) / GaussianBlur throw in the reference (:347-356) */
        throw AbubError("bellows template pasted outside the frame in event " + EventID + " camera " + std::to_string(CameraNumber));
    again.resize(n);
    return true;
}

void L3Localizer::TrackFromBlobs(const abub_blob* blobs, int n) {
    std::vector<BubbleImageFrame> cands;
    for (int i = 0; i < n; i++) {
        if (blobs[i].w * blobs[i].h > 10) {                                // :812
            BubbleImageFrame f = FrameFromBlob(blobs[i], false);
            if (!isInMask(&f.newPosition)) continue;                       // :826
            cands.push_back(f);
        }
    }
    /* :842-863: unlock every bubble, then each candidate (contour order) goes to the first bubble with
     * last_x - x < 5 (no abs) and |last_y - y| < 5; operator<< appends only to an unlocked bubble but the
     * candidate is consumed either way */
    for (bubble* b : BubbleList) b->lockThisIteration = false;
    for (const BubbleImageFrame& c : cands) {
        const float tx = c.MassCentres.x, ty = c.MassCentres.y;
        for (bubble* b : BubbleList) {
            if ((b->last_x - tx) < 5 && fabs(b->last_y - ty) < 5) {
                *b << c;
                break;
            }
        }
    }
}

void L3Localizer::LocalizeOMatic(std::string) {
    if (CameraFrames.size() <= 5) okToProceed = false;                     // :889
    if (!okToProceed) return;
    const int N = (int)CameraFrames.size();
    int prev_offset = 2;
    if (TrainedData->TrainingSetSize < 6) {                                // :919-926
        prev_offset = 1;
        if (!nonStopMode) std::cout << "Setting prev_offset to " << prev_offset << std::endl;
    }
    int preTrigNum = MatTrigFrame - prev_offset;
    if (preTrigNum < 0) preTrigNum = 0;
    preTrigNumber = preTrigNum;
    /* :944-954 */
    const int kmax = MatTrigFrame < 29 ? NumFramesBubbleTrack : 39 - MatTrigFrame;
    int n_post = 0;
    for (int k = 1; k <= kmax; k++) {
        if (MatTrigFrame + k >= N) break;
        n_post++;
    }
    EnsureFramesOnDevice();
    /* frames the reference decodes here: trigger, pre-trigger, frame 0 (:929-939) and trig+1..trig+n_post (:770);
     * an undecodable one ends in an OpenCV exception there (-6 in AnyCamAnalysis) */
    if (FrameIsBad(MatTrigFrame, MatTrigFrame + n_post) || FrameIsBad(preTrigNum, preTrigNum) || FrameIsBad(0, 0))
        throw AbubError("undecodable frame in event " + EventID + " camera " + std::to_string(CameraNumber));

    AbubDevice& dev = AbubDevice::Get();
    std::vector<abub_blob> blobs(1024);
    int npf[ABUB_MAX_TRACK + 1] = {0};
    {
        AbubSlot& st = dev.slot(Slot);
        const int c = CameraNumber;
        if (st.loc_valid[c] && st.loc_trig[c] == MatTrigFrame && st.loc_pre[c] == preTrigNum && st.loc_thres[c] == loc_thres &&
            st.loc_npost[c] == n_post) {                                    // localized with the rest of the batch
            blobs = st.loc_blobs[c];
            for (int k = 0; k <= ABUB_MAX_TRACK; k++) npf[k] = st.loc_npf[c][k];
        } else {
            std::lock_guard<std::recursive_mutex> g(dev.lock());
            dev.EnsureAllResident(Slot);
            for (;;) {
                const int rc = abub_localize(dev.ctx(), Slot, CameraNumber, MatTrigFrame, preTrigNum, loc_thres, n_post,
                                             blobs.data(), (int)blobs.size(), npf);
                if (rc == ABUB_E_CAPACITY && blobs.size() < (1u << 22)) { blobs.resize(blobs.size() * 4); continue; }
                dev.Check(rc, "abub_localize");
                break;
            }
        }
    }
    const abub_blob* p = blobs.data();
    GenesisFromBlobs(p, npf[0]);                                           // CalculateInitialBubbleParams
    p += npf[0];
    for (int k = 1; k <= n_post; k++) {                                    // CalculatePostTriggerFrameParams(k)
        TrackFromBlobs(p, npf[k]);
        p += npf[k];
    }
}
