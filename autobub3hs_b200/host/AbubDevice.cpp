#include "AbubDevice.hpp"

#include <string.h>

#include <algorithm>
#include <chrono>

#include "ImageCodec.hpp"

AbubDevice& AbubDevice::Get() {
    static AbubDevice dev;
    return dev;
}

void AbubDevice::Configure(int device, int ncams, int max_frames, int max_slots) {
    std::lock_guard<std::recursive_mutex> g(mu_);
    if (ctx_) throw AbubError("AbubDevice::Configure called after the context was created");
    if (ncams < 1 || ncams > 8 || max_frames < 2 || max_slots < 1) throw AbubError("AbubDevice::Configure: bad capacities");
    device_ = device; ncams_ = ncams; max_frames_ = max_frames; max_slots_ = max_slots;
    slots_.assign(max_slots, AbubSlot());
    by_key_.clear();
}

void AbubDevice::Check(int rc, const char* what) {
    if (rc != ABUB_OK) throw AbubError(std::string(what) + " failed (" + std::to_string(rc) + "): " + abub_last_error(ctx_));
}

abub_ctx* AbubDevice::Context(int rows, int cols) {
    std::lock_guard<std::recursive_mutex> g(mu_);
    if (!configured()) throw AbubError("AbubDevice is not configured");
    if (ctx_) {
        if (rows != rows_ || cols != cols_)
            throw AbubError("frame of " + std::to_string(cols) + "x" + std::to_string(rows) + " in a run of " +
                            std::to_string(cols_) + "x" + std::to_string(rows_));
        return ctx_;
    }
    abub_ctx* c = nullptr;
    const int rc = abub_ctx_create(device_, rows, cols, ncams_, max_frames_, max_slots_, &c);
    if (rc != ABUB_OK) {
        const std::string why = c ? abub_last_error(c) : "no CUDA device";
        if (c) abub_ctx_destroy(c);
        throw AbubError("abub_ctx_create failed (" + std::to_string(rc) + "): " + why +
                        " -- this build has no CPU path");
    }
    ctx_ = c; rows_ = rows; cols_ = cols;
    if (trigger_metric_ != ABUB_METRIC_SIGNIFICANCE || trigger_threshold_ > 0.f)
        Check(abub_set_trigger_metric(ctx_, trigger_metric_, trigger_threshold_), "abub_set_trigger_metric");
    return ctx_;
}

bool AbubDevice::EnsureBellowsTemplate(int cam, const std::string& mask_dir) {
    std::lock_guard<std::recursive_mutex> g(mu_);
    const std::string key = mask_dir + "|" + std::to_string(cam);
    auto it = templates_.find(key);
    if (it != templates_.end()) return it->second;
    bool ok = false;
    std::vector<uint8_t> bytes;
    Image8 t;
    if (ctx_ && abubcodec::read_file(mask_dir + "/cam" + std::to_string(cam) + "_bellows_template.png", bytes) &&
        abubcodec::decode(bytes.data(), bytes.size(), t) && !t.empty()) {
        Check(abub_set_bellows_template(ctx_, cam, t.ptr(), t.rows, t.cols), "abub_set_bellows_template");
        ok = true;
    }
    templates_[key] = ok;
    return ok;
}

const Image8* AbubDevice::Mask(int cam, bool bellows, const std::string& mask_dir) {
    std::lock_guard<std::mutex> g(tab_mu_);
    const std::string key = mask_dir + "|" + std::to_string(cam) + (bellows ? "|b" : "|m");
    auto it = masks_.find(key);
    if (it == masks_.end()) {
        Image8 m;
        std::vector<uint8_t> bytes;
        if (abubcodec::read_file(mask_dir + "/cam" + std::to_string(cam) + (bellows ? "_bellows" : "") + "_mask.bmp", bytes))
            abubcodec::decode(bytes.data(), bytes.size(), m);
        it = masks_.emplace(key, std::move(m)).first;
    }
    return it->second.empty() ? nullptr : &it->second;
}

void AbubDevice::Shutdown() {
    std::lock_guard<std::recursive_mutex> g(mu_);
    templates_.clear();
    masks_.clear();
    if (ctx_) abub_ctx_destroy(ctx_);
    ctx_ = nullptr; rows_ = cols_ = 0;
    by_key_.clear();
    for (AbubSlot& s : slots_) s = AbubSlot();
}

int AbubDevice::AcquireSlot(const std::string& key) {
    std::lock_guard<std::mutex> g(tab_mu_);
    auto it = by_key_.find(key);
    if (it != by_key_.end()) { slots_[it->second].refs++; return it->second; }
    for (int s = 0; s < (int)slots_.size(); s++)
        if (slots_[s].refs == 0) {
            slots_[s] = AbubSlot();
            slots_[s].key = key;
            slots_[s].refs = 1;
            by_key_[key] = s;
            return s;
        }
    throw AbubError("no free event slot (max_slots = " + std::to_string(max_slots_) + ")");
}

int AbubDevice::BindSlot(const std::string& key, int slot) {
    std::lock_guard<std::mutex> g(tab_mu_);
    if (slot < 0 || slot >= (int)slots_.size()) throw AbubError("BindSlot: no such slot");
    auto it = by_key_.find(key);
    if (it != by_key_.end()) {
        if (it->second != slot) throw AbubError("BindSlot: event " + key + " is bound to another slot");
        slots_[slot].refs++;
        return slot;
    }
    if (slots_[slot].refs != 0) throw AbubError("BindSlot: slot " + std::to_string(slot) + " is in use");
    // a trigger search started ahead of time for the frames this event is about to claim survives the reset
    const bool pre = slots_[slot].trig_prelaunched;
    int pre_nf[8];
    for (int c = 0; c < 8; c++) pre_nf[c] = slots_[slot].pre_nframes[c];
    slots_[slot] = AbubSlot();
    slots_[slot].trig_prelaunched = pre;
    for (int c = 0; c < 8; c++) slots_[slot].pre_nframes[c] = pre_nf[c];
    slots_[slot].key = key;
    slots_[slot].refs = 1;
    by_key_[key] = slot;
    return slot;
}

bool AbubDevice::SlotInUse(int slot) {
    std::lock_guard<std::mutex> g(tab_mu_);
    return slot >= 0 && slot < (int)slots_.size() && slots_[slot].refs != 0;
}

void AbubDevice::ReleaseSlot(const std::string& key) {
    std::lock_guard<std::mutex> g(tab_mu_);
    auto it = by_key_.find(key);
    if (it == by_key_.end()) return;
    if (--slots_[it->second].refs <= 0) {
        slots_[it->second] = AbubSlot();
        by_key_.erase(it);
    }
}

void AbubDevice::LoadCamera(int slot, int cam, Parser* parser, const std::string& event_id,
                            const std::vector<std::string>& frames) {
    if (slots_[slot].loaded[cam]) return;          // staged with its batch: nothing to do, and no need to wait for the device lock
    std::lock_guard<std::recursive_mutex> g(mu_);
    AbubSlot& st = slots_[slot];
    if (st.loaded[cam]) return;
    const int n = (int)frames.size();
    if (n > max_frames_) throw AbubError("event " + event_id + " has " + std::to_string(n) + " frames for camera " +
                                         std::to_string(cam) + " (capacity " + std::to_string(max_frames_) + ")");
    auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t0 = now();
    std::vector<char> ok(n, 0);
    if (!ctx_) {
        // size the context from the first frame whose header can be read
        std::vector<uint8_t> enc;
        for (int i = 0; i < n && !ctx_; i++) {
            int r, c;
            if (parser->GetImageBytes(event_id, frames[i], enc) != parser->FailureCode() && !enc.empty() &&
                abubcodec::peek_size(enc.data(), enc.size(), &r, &c))
                Context(r, c);
        }
        if (!ctx_) throw AbubError("no decodable frame to size the device context from (event " + event_id + ")");
    }
    uint8_t* pinned = nullptr;
    Check(abub_pinned_acquire(ctx_, slot, &pinned), "abub_pinned_acquire");
    const size_t npix = (size_t)rows_ * cols_;
    uint8_t* base = pinned + (size_t)cam * max_frames_ * npix;
    const double t1 = now();
    // every core reads and decodes its own frames straight into the pinned staging buffer (RawParser: one file per
    // frame; ZipParser: positional reads on the shared descriptor)
#pragma omp parallel for schedule(dynamic, 1)
    for (int i = 0; i < n; i++) {
        bool good = false;
        try {                                   // nothing may leave the parallel region: a frame that throws is a bad frame
            std::vector<uint8_t> enc;
            good = parser->GetImageBytes(event_id, frames[i], enc) != parser->FailureCode() && !enc.empty();
            if (good) good = abubcodec::decode_into(enc.data(), enc.size(), base + (size_t)i * npix, rows_, cols_);
        } catch (...) { good = false; }
        if (!good) memset(base + (size_t)i * npix, 0, npix);
        ok[i] = good;
    }
    st.bad[cam].clear();
    for (int i = 0; i < n; i++)
        if (!ok[i]) st.bad[cam].push_back(i);
    const double t2 = now();
    Check(abub_event_submit_cam(ctx_, slot, cam, n), "abub_event_submit_cam");     // asynchronous: the next camera decodes meanwhile
    const double t3 = now();
    t_read += t1 - t0; t_decode += t2 - t1; t_upload += t3 - t2;      // t_read = context sizing + pinned allocation
    st.nframes[cam] = n;
    st.resident_upto[cam] = n;
    st.loaded[cam] = true;
}

void AbubDevice::StageBatch(Parser* parser, const std::vector<int>& slots, const std::vector<std::string>& events) {
    auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t0 = now();
    struct Item { int slot, cam, frame; const std::string* event; std::string name; char state; uint32_t zlen; };   // state: 0 bad, 1 host-decoded, 2 device PNG job, 3 BMP pixel array read as stored (rows to be turned over on the device)
    std::vector<Item> items;
    std::vector<std::vector<int>> ncam(slots.size(), std::vector<int>(ncams_, -1));      // frames per (event, cam); -1 = not staged here
    for (size_t e = 0; e < events.size(); e++)
        for (int cam = 0; cam < ncams_; cam++) {
            AbubSlot& st = slots_[slots[e]];
            if (st.loaded[cam]) continue;
            std::vector<std::string> frames;
            parser->ParseAndSortFramesInFolder(events[e], cam, frames);
            if ((int)frames.size() > max_frames_) continue;                // the analyzer reports the capacity problem (-6)
            ncam[e][cam] = (int)frames.size();
            for (int f = 0; f < (int)frames.size(); f++) items.push_back(Item{slots[e], cam, f, &events[e], frames[f], 0, 0});
        }
    if (!ctx_) {                                                           // size the context from the first readable header
        std::vector<uint8_t> enc;
        for (size_t i = 0; i < items.size() && !ctx_; i++) {
            int r, c;
            if (parser->GetImageBytes(*items[i].event, items[i].name, enc) != parser->FailureCode() && !enc.empty() &&
                abubcodec::peek_size(enc.data(), enc.size(), &r, &c))
                Context(r, c);
        }
        if (!ctx_) return;                                                 // nothing decodable: the analyzers report it
    }
    const size_t npix = (size_t)rows_ * cols_;
    std::map<int, uint8_t*> pin;
    for (int s : slots) {
        uint8_t* p = nullptr;
        Check(abub_pinned_acquire(ctx_, s, &p), "abub_pinned_acquire");
        pin[s] = p;
    }
    const double t1 = now();
    const bool device_png = (npix % 4 == 0) && !getenv("ABUB_PNG_ON_HOST");
    // 8-bit BMP frames (what the DAQ writes in raw mode): the pixel array goes from the file into the pinned region with
    // one positional read and the bottom-up row order is undone on the device after the upload
    const bool device_flip = (cols_ % 4 == 0) && !getenv("ABUB_BMP_ON_HOST");
    std::vector<abub_frame_ref> flips;
    // One event at a time: its frames are read and decoded on all cores, then the cameras that hold host-decoded frames go
    // up asynchronously while the next event is being decoded (their device-decoded frames carry compressed bytes for the
    // moment).  PNG streams are collected for one decode call over the whole batch (the inflate kernel wants many frames).
    std::vector<abub_png_job> jobs;
    std::vector<size_t> job_item;
    double t_dec = 0, t_up = 0;
    size_t i0 = 0;
    for (size_t e = 0; e < events.size(); e++) {
        size_t i1 = i0;
        while (i1 < items.size() && items[i1].event == &events[e]) i1++;
        const double ta = now();
#pragma omp parallel for schedule(dynamic, 1)
        for (long i = (long)i0; i < (long)i1; i++) {
            Item& it = items[i];
            uint8_t* region = pin[it.slot] + ((size_t)it.cam * max_frames_ + it.frame) * npix;
            try {                               // nothing may leave the parallel region: a frame that throws is a bad frame
            if (device_flip) {
                bool bottom_up = false;
                if (parser->ReadGray8BmpPixels(*it.event, it.name, rows_, cols_, region, &bottom_up)) { it.state = bottom_up ? 3 : 1; continue; }
            }
            static thread_local std::vector<uint8_t> enc;          // reused: no fresh 1.8 MB mapping per frame
            if (parser->GetImageBytes(*it.event, it.name, enc) == parser->FailureCode() || enc.empty()) { memset(region, 0, npix); continue; }
            if (device_png) {
                const size_t zl = abubcodec::png_gray8_stream(enc.data(), enc.size(), rows_, cols_, region, npix);
                if (zl >= 2) { it.state = 2; it.zlen = (uint32_t)zl; continue; }
            }
            if (abubcodec::decode_into(enc.data(), enc.size(), region, rows_, cols_)) it.state = 1;
            else memset(region, 0, npix);
            } catch (...) { it.state = 0; memset(region, 0, npix); }
        }
        const double tb = now();
        bool host_cam[8] = {false};
        for (size_t i = i0; i < i1; i++) {
            if (items[i].state == 2) { jobs.push_back(abub_png_job{items[i].slot, items[i].cam, items[i].frame, items[i].zlen}); job_item.push_back(i); }
            else host_cam[items[i].cam] = true;
        }
        const double ts0 = now();
        for (int cam = 0; cam < ncams_; cam++)
            if (ncam[e][cam] >= 0 && host_cam[cam]) Check(abub_event_submit_cam(ctx_, slots[e], cam, ncam[e][cam]), "abub_event_submit_cam");
        const double ts1 = now();
        t_submit += ts1 - ts0;
        flips.clear();
        for (size_t i = i0; i < i1; i++)
            if (items[i].state == 3) flips.push_back(abub_frame_ref{items[i].slot, items[i].cam, items[i].frame});
        if (!flips.empty()) Check(abub_event_flip_rows(ctx_, flips.data(), (int)flips.size()), "abub_event_flip_rows");   // after the uploads, same stream
        t_flip += now() - ts1;
        t_dec += tb - ta;
        t_up += now() - tb;
        i0 = i1;
    }
    const double t2 = t1 + t_dec;
    // ... then the PNG streams are sent and decoded in place
    std::vector<int> okv(jobs.size(), 0);
    if (!jobs.empty()) Check(abub_event_decode_png(ctx_, jobs.data(), (int)jobs.size(), okv.data()), "abub_event_decode_png");
    for (size_t j = 0; j < jobs.size(); j++) {
        Item& it = items[job_item[j]];
        if (okv[j]) { it.state = 1; n_gpu_decoded++; continue; }
        // the device rejected the stream: let the host decoder have the last word on this frame
        std::vector<uint8_t> enc;
        uint8_t* region = pin[it.slot] + ((size_t)it.cam * max_frames_ + it.frame) * npix;
        it.state = 0;
        if (parser->GetImageBytes(*it.event, it.name, enc) != parser->FailureCode() && !enc.empty() &&
            abubcodec::decode_into(enc.data(), enc.size(), region, rows_, cols_)) {
            it.state = 1;
            Check(abub_event_upload_frame(ctx_, it.slot, it.cam, it.frame), "abub_event_upload_frame");
        }
    }
    for (size_t e = 0; e < events.size(); e++) {
        AbubSlot& st = slots_[slots[e]];
        int nf[8] = {0};
        for (int cam = 0; cam < ncams_; cam++) {
            if (ncam[e][cam] >= 0) {
                st.nframes[cam] = ncam[e][cam];
                st.resident_upto[cam] = ncam[e][cam];
                st.loaded[cam] = true;
                st.bad[cam].clear();
            }
            nf[cam] = st.nframes[cam];
        }
        Check(abub_event_set_nframes(ctx_, slots[e], nf), "abub_event_set_nframes");
    }
    for (const Item& it : items) {
        if (it.state == 0) slots_[it.slot].bad[it.cam].push_back(it.frame);
        else if (it.zlen == 0) n_cpu_decoded++;
    }
    const double t3 = now();
    t_read += t1 - t0; t_decode += t2 - t1; t_upload += t3 - t2;
    t_tail += (t3 - t2) - t_up;
}

std::vector<int> AbubDevice::PrefetchEvents(Parser* parser, const std::vector<std::string>& keys, const std::vector<std::string>& events,
                                            const int* training_set_sizes) {
    std::lock_guard<std::recursive_mutex> g(mu_);
    std::vector<int> out;
    for (size_t e = 0; e < events.size(); e++) out.push_back(AcquireSlot(keys[e]));
    if (getenv("ABUB_HOST_DECODE")) {
        for (size_t e = 0; e < events.size(); e++)
            for (int cam = 0; cam < ncams_; cam++) {
                std::vector<std::string> frames;
                parser->ParseAndSortFramesInFolder(events[e], cam, frames);
                try {
                    LoadCamera(out[e], cam, parser, events[e], frames);
                } catch (const AbubError&) {
                    // left unloaded: the analyzer of this camera will hit the same error and report -6
                }
            }
    } else {
        StageBatch(parser, out, events);
    }
    AnalyzeBatch(out, training_set_sizes);
    return out;
}

static inline double wall_now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

void AbubDevice::MarkResident(int slot, const int* nframes, bool in_hbm, bool push) {
    std::lock_guard<std::recursive_mutex> g(mu_);
    AbubSlot& st = slots_[slot];
    for (int c = 0; c < ncams_; c++) {
        st.nframes[c] = nframes[c];
        st.loaded[c] = true;
        st.bad[c].clear();
        st.resident_upto[c] = in_hbm ? nframes[c] : 0;
        st.trig_valid[c] = st.loc_valid[c] = false;
    }
    if (!in_hbm) st.trig_prelaunched = false;        // the frames come from the staging buffer: nothing started ahead counts
    if (in_hbm && push) Check(abub_event_set_nframes(ctx_, slot, nframes), "abub_event_set_nframes");
}

// consecutive slots whose frames are in HBM: the device's frame counts in one launch (after MarkResident(..., true, false))
void AbubDevice::PushFrameCounts(int slot_first, int nslots, const int* nframes) {
    std::lock_guard<std::recursive_mutex> g(mu_);
    Check(abub_event_set_nframes_range(ctx_, slot_first, nslots, nframes), "abub_event_set_nframes_range");
}

void AbubDevice::EnsureAllResident(int slot) {
    std::lock_guard<std::recursive_mutex> g(mu_);
    AbubSlot& st = slots_[slot];
    for (int c = 0; c < ncams_; c++)
        if (st.loaded[c] && st.resident_upto[c] < st.nframes[c]) {
            Check(abub_event_submit_range(ctx_, slot, c, st.resident_upto[c], st.nframes[c] - st.resident_upto[c]), "abub_event_submit_range");
            st.resident_upto[c] = st.nframes[c];
        }
}

// FindTriggerFrame(startframe = 1) of every (event, camera) of the slots, then LocalizeOMatic of every triggered one.
// Registers the early-exit trigger search of slots [slot_first, slot_first + nslots) -- frames resident, `nframes` per camera --
// to be launched inside the next localizer call.  False (nothing registered) when the full-table form is configured or a slot
// is still held by an analyzer.
bool AbubDevice::PrefetchTrigger(int slot_first, int nslots, const int* nframes) {
    std::lock_guard<std::recursive_mutex> g(mu_);
    if (!ctx_ || early_chunk_ <= 0 || slot_first < 0 || nslots < 1 || slot_first + nslots > max_slots_) return false;
    {
        std::lock_guard<std::mutex> t(tab_mu_);
        for (int s = slot_first; s < slot_first + nslots; s++)
            if (slots_[s].refs > 0) return false;
    }
    Check(abub_event_set_nframes_range(ctx_, slot_first, nslots, nframes), "abub_event_set_nframes_range");
    Check(abub_trigger_early_deferred(ctx_, slot_first, nslots, early_chunk_), "abub_trigger_early_deferred");
    for (int s = 0; s < nslots; s++) {
        AbubSlot& st = slots_[slot_first + s];
        st.trig_prelaunched = true;
        for (int c = 0; c < ncams_; c++) st.pre_nframes[c] = nframes[(size_t)s * ncams_ + c];
    }
    return true;
}

void AbubDevice::AnalyzeBatch(const std::vector<int>& out, const int* training_set_sizes, int next_first, int next_n,
                              const int* next_nframes) {
    std::lock_guard<std::recursive_mutex> g(mu_);
    if (!ctx_) return;
    std::vector<int> sorted(out);
    std::sort(sorted.begin(), sorted.end());
    const double h0 = wall_now();
    // searches started ahead of time (PrefetchTrigger) must be launched in full before their results are read
    bool any_pre = false;
    for (int s : sorted) any_pre = any_pre || slots_[s].trig_prelaunched;
    if (any_pre) Check(abub_trigger_deferred_flush(ctx_), "abub_trigger_deferred_flush");
    // The trigger search of a contiguous slot range.  Default: the resident early exit (abub_trigger_early) -- K1 over
    // early_chunk_ frame pairs at a time, K2 after every round, decided cameras drop out, as the reference stops calling
    // ProcessFrame at the trigger (AnalyzerUnit.cpp:297-300).  early_chunk_ = 0: the full table (one K1 launch over all
    // pairs, then K2).  A re-trigger (startframe > 1) rebuilds the slot's table on demand (abub_trigger).
    for (size_t i = 0; i < sorted.size();) {
        size_t j = i + 1;
        while (j < sorted.size() && sorted[j] == sorted[j - 1] + 1) j++;
        const int first = sorted[i], ns = (int)(j - i);
        bool pre = true;
        for (int s = first; s < first + ns; s++) {
            pre = pre && slots_[s].trig_prelaunched;
            for (int c = 0; c < ncams_ && pre; c++) pre = slots_[s].pre_nframes[c] == slots_[s].nframes[c];
        }
        for (int s = first; s < first + ns; s++) slots_[s].trig_prelaunched = false;
        if (pre) {
            // started inside the previous batch's localizer call: nothing to launch
        } else if (early_chunk_ > 0) {
            Check(abub_trigger_early(ctx_, first, ns, early_chunk_), "abub_trigger_early");
        } else {
            Check(abub_trigger_hist(ctx_, first, ns), "abub_trigger_hist");
            Check(abub_trigger_scan(ctx_, first, ns, nullptr), "abub_trigger_scan");
        }
        i = j;
    }
    const double h1 = wall_now();
    t_hist += h1 - h0;
    for (size_t i = 0; i < sorted.size();) {
        size_t j = i + 1;
        while (j < sorted.size() && sorted[j] == sorted[j - 1] + 1) j++;
        const int first = sorted[i], ns = (int)(j - i);
        std::vector<int> tr(ns * ncams_), stt(ns * ncams_), lt(ns * ncams_);
        std::vector<double> sg((size_t)ns * ncams_ * max_frames_);
        Check(abub_trigger_results(ctx_, first, ns, tr.data(), stt.data(), lt.data(), sg.data()), "abub_trigger_results");
        for (int s = 0; s < ns; s++)
            for (int c = 0; c < ncams_; c++) {
                AbubSlot& st = slots_[first + s];
                if (!st.loaded[c] || st.nframes[c] < 5) continue;
                const int k = s * ncams_ + c;
                st.trig[c] = tr[k]; st.trig_status[c] = stt[k]; st.trig_loc_thres[c] = lt[k];
                st.sig[c].assign(sg.begin() + (size_t)k * max_frames_, sg.begin() + (size_t)(k + 1) * max_frames_);
                st.trig_valid[c] = true;
            }
        i = j;
    }
    const double h2 = wall_now();
    t_scan += h2 - h1;
    // the next batch's search: registered, its first round launched now (device work while the item and job tables of this
    // batch's localizer call are built), the rest by the localizer call itself
    if (next_first >= 0 && next_n > 0 && next_nframes && PrefetchTrigger(next_first, next_n, next_nframes))
        Check(abub_trigger_deferred_start(ctx_), "abub_trigger_deferred_start");
    if (training_set_sizes) LocalizeBatch(out, training_set_sizes);
    Check(abub_trigger_deferred_flush(ctx_), "abub_trigger_deferred_flush");     // no-op when the localizer call launched it
    t_loc += wall_now() - h2;
}

// Events whose frames wait in the slots' pinned staging buffers (MarkResident(..., false)): abub_trigger_stream sends them
// chunk by chunk and stops per camera at the decision; what the localizer still needs (up to trig + n_post) follows.
long long AbubDevice::StreamBatch(const std::vector<int>& out, const int* training_set_sizes, int chunk_frames) {
    std::lock_guard<std::recursive_mutex> g(mu_);
    if (!ctx_) return 0;
    std::vector<int> sorted(out);
    std::sort(sorted.begin(), sorted.end());
    long long frames_sent = 0;
    const double h0 = wall_now();
    for (size_t i = 0; i < sorted.size();) {
        size_t j = i + 1;
        while (j < sorted.size() && sorted[j] == sorted[j - 1] + 1) j++;
        const int first = sorted[i], ns = (int)(j - i);
        std::vector<int> nf(ns * ncams_), tr(ns * ncams_), stt(ns * ncams_), lt(ns * ncams_), sent(ns * ncams_);
        for (int s = 0; s < ns; s++)
            for (int c = 0; c < ncams_; c++) nf[s * ncams_ + c] = slots_[first + s].nframes[c];
        Check(abub_trigger_stream(ctx_, first, ns, nf.data(), chunk_frames, tr.data(), stt.data(), lt.data(), sent.data()), "abub_trigger_stream");
        for (int s = 0; s < ns; s++)
            for (int c = 0; c < ncams_; c++) {
                AbubSlot& st = slots_[first + s];
                const int k = s * ncams_ + c;
                st.resident_upto[c] = sent[k];
                frames_sent += sent[k];
                if (!st.loaded[c] || st.nframes[c] < 5) continue;
                st.trig[c] = tr[k]; st.trig_status[c] = stt[k]; st.trig_loc_thres[c] = lt[k];
                st.sig[c].assign(max_frames_, 0.0);          // the streamed search does not read the per-frame values back
                st.trig_valid[c] = true;
            }
        i = j;
    }
    const double h1 = wall_now();
    t_hist += h1 - h0;
    if (training_set_sizes) {
        // frames trig+1 .. trig+n_post of every triggered camera (L3Localizer.cpp:770) that the search did not need
        for (int s : out)
            for (int c = 0; c < ncams_; c++) {
                AbubSlot& st = slots_[s];
                if (!st.trig_valid[c] || st.trig_status[c] != 0 || st.nframes[c] <= 5) continue;
                const int need = st.trig[c] + NPost(st.trig[c], st.nframes[c]) + 1;
                if (need > st.resident_upto[c]) {
                    Check(abub_event_submit_range(ctx_, s, c, st.resident_upto[c], need - st.resident_upto[c]), "abub_event_submit_range");
                    frames_sent += need - st.resident_upto[c];
                    st.resident_upto[c] = need;
                }
            }
        LocalizeBatch(out, training_set_sizes);
    }
    t_loc += wall_now() - h1;
    return frames_sent;
}

int AbubDevice::NPost(int trig, int N) {                         // L3Localizer.cpp:944-954
    const int kmax = trig < 29 ? ABUB_MAX_TRACK : 39 - trig;
    int n_post = 0;
    for (int k = 1; k <= kmax; k++) {
        if (trig + k >= N) break;
        n_post++;
    }
    return n_post;
}

// LocalizeOMatic of every triggered (event, camera): one abub_localize_batch
void AbubDevice::LocalizeBatch(const std::vector<int>& out, const int* training_set_sizes) {
    std::vector<abub_loc_item> items;
    std::vector<std::pair<int, int>> who;
    for (int s : out)
        for (int c = 0; c < ncams_; c++) {
            AbubSlot& st = slots_[s];
            if (!st.trig_valid[c] || st.trig_status[c] != 0 || st.nframes[c] <= 5 || !st.bad[c].empty()) continue;
            const int trig = st.trig[c], N = st.nframes[c];
            int pre = trig - (training_set_sizes[c] < 6 ? 1 : 2);                    // L3Localizer.cpp:919-934
            if (pre < 0) pre = 0;
            items.push_back(abub_loc_item{s, c, trig, pre, st.trig_loc_thres[c], NPost(trig, N)});
            who.push_back({s, c});
        }
    if (items.empty()) return;
    if (loc_blobs_.size() < ((size_t)1 << 16)) loc_blobs_.resize((size_t)1 << 16);
    std::vector<int> npf(items.size() * (ABUB_MAX_TRACK + 1));
    int total = 0;
    for (;;) {
        const int rc = abub_localize_batch(ctx_, items.data(), (int)items.size(), loc_blobs_.data(), (int)loc_blobs_.size(), npf.data(), &total);
        if (rc == ABUB_E_CAPACITY && loc_blobs_.size() < ((size_t)1 << 26)) { loc_blobs_.resize((size_t)total + 1024); continue; }
        Check(rc, "abub_localize_batch");
        break;
    }
    size_t off = 0;
    for (size_t it = 0; it < items.size(); it++) {
        AbubSlot& st = slots_[who[it].first];
        const int c = who[it].second;
        size_t cnt = 0;
        for (int k = 0; k <= ABUB_MAX_TRACK; k++) {
            st.loc_npf[c][k] = npf[it * (ABUB_MAX_TRACK + 1) + k];
            cnt += (size_t)st.loc_npf[c][k];
        }
        st.loc_blobs[c].assign(loc_blobs_.begin() + off, loc_blobs_.begin() + off + cnt);
        off += cnt;
        st.loc_trig[c] = items[it].trig; st.loc_pre[c] = items[it].pre_trig; st.loc_thres[c] = items[it].loc_thres;
        st.loc_npost[c] = items[it].n_post;
        st.loc_valid[c] = true;
    }
}

