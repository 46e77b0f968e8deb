#include "AbubDevice.hpp"

#include <string.h>

#include <algorithm>

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

void AbubDevice::Shutdown() {
    std::lock_guard<std::recursive_mutex> g(mu_);
    templates_.clear();
    if (ctx_) abub_ctx_destroy(ctx_);
    ctx_ = nullptr; rows_ = cols_ = 0;
    by_key_.clear();
    for (AbubSlot& s : slots_) s = AbubSlot();
}

int AbubDevice::AcquireSlot(const std::string& key) {
    std::lock_guard<std::recursive_mutex> g(mu_);
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

void AbubDevice::ReleaseSlot(const std::string& key) {
    std::lock_guard<std::recursive_mutex> g(mu_);
    auto it = by_key_.find(key);
    if (it == by_key_.end()) return;
    if (--slots_[it->second].refs <= 0) {
        slots_[it->second] = AbubSlot();
        by_key_.erase(it);
    }
}

void AbubDevice::LoadCamera(int slot, int cam, Parser* parser, const std::string& event_id,
                            const std::vector<std::string>& frames) {
    std::lock_guard<std::recursive_mutex> g(mu_);
    AbubSlot& st = slots_[slot];
    if (st.loaded[cam]) return;
    const int n = (int)frames.size();
    if (n > max_frames_) throw AbubError("event " + event_id + " has " + std::to_string(n) + " frames for camera " +
                                         std::to_string(cam) + " (capacity " + std::to_string(max_frames_) + ")");
    // the encoded bytes are read serially (a parser owns one file handle), decoded on every core
    std::vector<std::vector<uint8_t>> enc(n);
    std::vector<char> ok(n, 0);
    for (int i = 0; i < n; i++) ok[i] = parser->GetImageBytes(event_id, frames[i], enc[i]) != parser->FailureCode() && !enc[i].empty();
    if (!ctx_) {
        for (int i = 0; i < n && !ctx_; i++) {
            int r, c;
            if (ok[i] && abubcodec::peek_size(enc[i].data(), enc[i].size(), &r, &c)) Context(r, c);
        }
        if (!ctx_) throw AbubError("no decodable frame to size the device context from (event " + event_id + ")");
    }
    uint8_t* pinned = nullptr;
    Check(abub_pinned_acquire(ctx_, slot, &pinned), "abub_pinned_acquire");
    const size_t npix = (size_t)rows_ * cols_;
    uint8_t* base = pinned + (size_t)cam * max_frames_ * npix;
#pragma omp parallel for schedule(dynamic, 1)
    for (int i = 0; i < n; i++) {
        if (ok[i]) ok[i] = abubcodec::decode_into(enc[i].data(), enc[i].size(), base + (size_t)i * npix, rows_, cols_);
        if (!ok[i]) memset(base + (size_t)i * npix, 0, npix);
    }
    st.bad[cam].clear();
    for (int i = 0; i < n; i++)
        if (!ok[i]) st.bad[cam].push_back(i);
    Check(abub_event_upload(ctx_, slot, cam, base, n), "abub_event_upload");
    st.nframes[cam] = n;
    st.loaded[cam] = true;
}

std::vector<int> AbubDevice::PrefetchEvents(Parser* parser, const std::vector<std::string>& keys, const std::vector<std::string>& events) {
    std::lock_guard<std::recursive_mutex> g(mu_);
    std::vector<int> out;
    for (size_t e = 0; e < events.size(); e++) {
        const std::string& ev = events[e];
        const int s = AcquireSlot(keys[e]);
        out.push_back(s);
        for (int cam = 0; cam < ncams_; cam++) {
            std::vector<std::string> frames;
            parser->ParseAndSortFramesInFolder(ev, cam, frames);
            try {
                LoadCamera(s, cam, parser, ev, frames);
            } catch (const AbubError&) {
                // left unloaded: the analyzer of this camera will hit the same error and report -6
            }
        }
    }
    // one K1 launch per contiguous slot range
    std::vector<int> sorted(out);
    std::sort(sorted.begin(), sorted.end());
    size_t i = 0;
    while (ctx_ && i < sorted.size()) {
        size_t j = i + 1;
        while (j < sorted.size() && sorted[j] == sorted[j - 1] + 1) j++;
        Check(abub_trigger_hist(ctx_, sorted[i], (int)(j - i)), "abub_trigger_hist");
        i = j;
    }
    return out;
}
