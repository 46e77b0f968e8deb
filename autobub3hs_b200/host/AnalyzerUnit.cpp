#include "AnalyzerUnit.hpp"

#include <iostream>

#include "AbubDevice.hpp"

AnalyzerUnit::AnalyzerUnit(std::string EventID_, std::string ImageDir_, int CameraNumber_, Trainer** TrainedData_,
                           std::string MaskDir_, Parser* parser) {
    ImageDir = ImageDir_;
    MaskDir = MaskDir_;
    CameraNumber = CameraNumber_;
    EventID = EventID_;
    TrainedData = *TrainedData_;
    MatTrigFrame = 0;
    loc_thres = 3;
    FileParser = parser;
    FileParser->ParseAndSortFramesInFolder(EventID, CameraNumber, CameraFrames);      // AnalyzerUnit.cpp:33
    SlotKey = ImageDir + "|" + EventID;
}

AnalyzerUnit::~AnalyzerUnit(void) {
    for (bubble* b : BubbleList) delete b;
    if (Slot >= 0) AbubDevice::Get().ReleaseSlot(SlotKey);
    if (FileParser) delete FileParser;
}

void AnalyzerUnit::EnsureFramesOnDevice() {
    AbubDevice& dev = AbubDevice::Get();
    TrainedData->InstallOnDevice();
    if (Slot < 0) Slot = dev.AcquireSlot(SlotKey);
    dev.LoadCamera(Slot, CameraNumber, FileParser, EventID, CameraFrames);
}

bool AnalyzerUnit::FrameIsBad(int first, int last) const {
    for (int b : AbubDevice::Get().slot(Slot).bad[CameraNumber])
        if (b >= first && b <= last) return true;
    return false;
}

void AnalyzerUnit::FindTriggerFrame(bool nonStopMode, int startframe) {
    /* malformed sequence (AnalyzerUnit.cpp:122-126) */
    if (CameraFrames.size() < 5) {
        okToProceed = false;
        TriggerFrameIdentificationStatus = -9;
        return;
    }
    if (startframe < 1) startframe = 1;
    if (startframe == 1) entropies.clear();
    const int N = (int)CameraFrames.size();
    AbubDevice& dev = AbubDevice::Get();
    EnsureFramesOnDevice();
    if (!nonStopMode)
        std::cout << "twoFrameOffset: " << (TrainedData->TrainingSetSize >= 6)
                  << "; this->TrainedData->TrainingSetSize: " << TrainedData->TrainingSetSize << std::endl;

    int trig = 0, status = -3, lt = -1;
    std::vector<double> sig(dev.max_frames(), 0.0);
    {
        AbubSlot& st = dev.slot(Slot);
        if (startframe == 1 && st.trig_valid[CameraNumber]) {      // scanned with the rest of the batch (AnalyzeBatch / StreamBatch)
            trig = st.trig[CameraNumber]; status = st.trig_status[CameraNumber]; lt = st.trig_loc_thres[CameraNumber];
            sig = st.sig[CameraNumber];
        } else {
            std::lock_guard<std::recursive_mutex> g(dev.lock());
            dev.EnsureAllResident(Slot);                  // a streamed batch stopped sending frames at the first decision
            dev.Check(abub_trigger(dev.ctx(), Slot, CameraNumber, startframe, &trig, &status, &lt, sig.data()), "abub_trigger");
        }
    }
    /* frames the reference would have read: startframe-2 .. trig+2 on a trigger, everything otherwise.  An
     * unreadable frame among them stops the search: -9 through the zip reader (AnalyzerUnit.cpp:207-213);
     * through the raw reader the empty cv::Mat makes OpenCV throw, which AnyCamAnalysis reports as -6. */
    const int last_read = (status == 0) ? (trig + 2 < N - 1 ? trig + 2 : N - 1) : N - 1;
    if (FrameIsBad(startframe >= 2 ? startframe - 2 : 0, last_read)) {
        if (FileParser->FailureCode() == -1) {
            std::cout << "Image of camera " << CameraNumber << " for the event " << EventID << " is corrupted/empty." << std::endl;
            okToProceed = false;
            TriggerFrameIdentificationStatus = -9;
            return;
        }
        throw AbubError("undecodable frame in event " + EventID + " camera " + std::to_string(CameraNumber));
    }
    const int last_stored = (status == 0) ? trig : N - 1;
    for (int i = startframe; i <= last_stored; i++) entropies.push_back(sig[i]);
    if (lt >= 0) loc_thres = lt;
    TriggerFrameIdentificationStatus = status;
    if (status == 0) {
        MatTrigFrame = trig;
    } else {
        okToProceed = false;                                     // AnalyzerUnit.cpp:317-321
    }
}
