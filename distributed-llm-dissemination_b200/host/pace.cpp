#include "pace.h"

#include <algorithm>

namespace dissem {
namespace host {

std::vector<PacePiece> PaceSchedule(int64_t layer_off, int64_t size, int64_t rate) {
    std::vector<PacePiece> out;
    if (size <= 0) return out;
    if (rate <= 0) {
        out.push_back(PacePiece{layer_off, size, 0});
        return out;
    }
    // merge sub-writes so that launches are at least ~250 us apart: quantum = a multiple of the burst
    const __int128 per_quarter_ms = (__int128)rate / 4000;
    int64_t quantum = kPaceBurst;
    if (per_quarter_ms > kPaceBurst) {
        const __int128 q = (per_quarter_ms + kPaceBurst - 1) / kPaceBurst * kPaceBurst;
        quantum = q > ((__int128)1 << 40) ? ((int64_t)1 << 40) : (int64_t)q;
    }
    const int64_t end = layer_off + size;
    for (int64_t pos = layer_off; pos < end;) {
        const int64_t piece_end = std::min(end, (pos / quantum + 1) * quantum);
        // The reference cuts its sub-writes from the FRAGMENT's start (data[:n]; data = data[n:]).  This piece may be
        // launched when the sub-write that holds its last byte is due: that one ends at cumulative byte sub_end.
        const int64_t cum_end = piece_end - layer_off;
        const int64_t sub_end = std::min(size, (cum_end + kPaceBurst - 1) / kPaceBurst * kPaceBurst);
        const int64_t due = std::max<int64_t>(0, sub_end - kPaceBurst);
        PacePiece p;
        p.off = pos;
        p.size = piece_end - pos;
        p.release_ns = (int64_t)((__int128)due * 1000000000 / rate);
        out.push_back(p);
        pos = piece_end;
    }
    return out;
}

}  // namespace host
}  // namespace dissem
