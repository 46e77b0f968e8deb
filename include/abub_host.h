/* abub_host.h -- entry points of libabubhost.so (autobub3hs_b200/host/ReconDriver.cpp).
 *
 * libabubhost.so holds the host C++ side of the drop-in -- Trainer / AnalyzerUnit / L3Localizer / bubble /
 * OutputWriter with the reference's class surface (AnalyzerUnit.hpp:20-100, L3Localizer.hpp:27-108,
 * Trainer.hpp:18-75, PICOFormatWriterV4.hpp:15-62) written against include/abub_cuda.h -- and the detect loop of
 * main (AutoBubStart3.cpp:342-388: events in parallel, cameras in series through AnyCamAnalysis :67-125, rows
 * appended per event inside `omp ordered` :379-383).  The abub3hs binary drives that loop from files; the calls
 * below drive the SAME loop for events whose frames a producer has placed in memory: in the HBM slots of the
 * context (abub_event_device_ptr) or in their pinned staging buffers (abub_pinned_acquire).  bench.py and the
 * GPU tests use them; a DAQ process that holds frames in RAM would too.
 *
 * int return: 0 = ok, -1 = failed (abubhost_last_error()).  One run at a time per process (AbubDevice is a singleton,
 * as the reference's main owns one OutputWriter and one set of Trainers).
 */
#ifndef ABUB_HOST_H
#define ABUB_HOST_H

#include "abub_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

const char* abubhost_last_error(void);

/* AbubDevice::Configure + context creation.  early_chunk = frame pairs per round of the resident early exit of the
 * trigger search (abub_trigger_early; AnalyzerUnit.cpp:297-300), 0 = all frame pairs in one K1 launch.  *ctx = the
 * abub_ctx of include/abub_cuda.h (owned by the library; do not destroy it). */
int abubhost_open(int device, int rows, int cols, int ncams, int max_frames, int max_slots, int early_chunk, abub_ctx** ctx);

/* Replaces main's set-up before the event loop (AutoBubStart3.cpp:250-251 writeHeader, :295-298 Trainers): the header
 * is written to <out_dir>/abub3hs_<run_id>.txt (truncated) and one Trainer per camera adopts the planes the caller has
 * installed on the device (abub_train_* + abub_set_background_from_training, or abub_set_background);
 * training_set_sizes[ncams] = Trainer::TrainingSetSize.  mask_dir = the -c argument ("" = no mask check). */
int abubhost_begin_run(const char* run_id, const char* out_dir, const char* mask_dir, int frame_offset,
                       const int* training_set_sizes);

/* One batch of the event loop (AutoBubStart3.cpp:342-388).  Slots [slot_first, slot_first+nslots) hold the events
 * event_numbers[nslots] (the `actualEventNumber` of the output rows) with nframes[nslots*ncams] frames per camera.
 * mode 0: the frames are in HBM.  mode 1: they wait in the pinned staging buffers and cross PCIe chunk_frames at a
 * time, per camera only as far as FindTriggerFrame reads (trig+2, AnalyzerUnit.cpp:191-300) and
 * CalculatePostTriggerFrameParams needs (trig+n_post, L3Localizer.cpp:770).
 * Per event: FindTriggerFrame -> LocalizeOMatic -> stageCameraOutput -> rows appended and flushed, in event order.
 * stats[8] (may be NULL): stats[5] is incremented by the frames sent over PCIe (the other entries are reserved; stage
 * times come from abubhost_stats). */
int abubhost_recon(int slot_first, int nslots, const int* event_numbers, const int* nframes, int mode, int chunk_frames,
                   double* stats);

/* Optional.  Announces the NEXT abubhost_recon call (mode 0): exactly these slots, these frame counts, frames already in
 * HBM.  The current call then queues that batch's trigger search inside its own localizer call, so that the GPU has work
 * during the localizer's host round trips and the bookkeeping between two calls (a caller that alternates between two sets
 * of slots announces the other set before each call).  A hint that does not come true costs one trigger search. */
int abubhost_recon_hint(int slot_first, int nslots, const int* nframes);

/* abubhost_recon returns once the device work of its events is done; the per-camera objects of the last sub-batch may
 * still be running on the host cores (they overlap the device work of the next call).  abubhost_drain returns when
 * every submitted event has been analysed and its rows are in the file. */
int abubhost_drain(void);
/* Host threads of the event loop (0 = OpenMP's default; a launcher such as torchrun exports OMP_NUM_THREADS=1) and events
 * per sub-batch, the unit handed from the device stage to the host stage (0 = a whole abubhost_recon call).  Take
 * effect at the next abubhost_begin_run. */
int abubhost_set_threads(int threads, int sub_batch);
/* Seconds per stage since the previous call (drains first): out[5] = {trigger launches, trigger read-back incl. waiting
 * for K1/K2, localizer, host analysis, file writes}. */
int abubhost_stats(double* out);

int abubhost_end_run(void);      /* drains and closes the output file */
void abubhost_close(void);       /* end_run + destroys the context */

#ifdef __cplusplus
}
#endif
#endif /* ABUB_HOST_H */
