// LimitRate pacing: the token bucket of TcpTransport.writeWithLimit (distributor/transport.go:407-424) and of the
// client branch of handleFlowRetransmit (node.go:1616-1626) — rate.NewLimiter(rate.Limit(limitRate), 256 KiB) from
// golang.org/x/time/rate v0.14.0 (go.mod:8; its source is not under /root/reference) — restated as a release schedule.
//
// The reference writes a fragment in sub-writes of at most burst = 256 KiB, each after limiter.WaitN(n).  The bucket
// starts full and the writer never idles, so the sub-write that ends at cumulative byte b is released at
//     max(0, b - burst) / rate   seconds after the first one
// (by then burst + rate*t >= b tokens have been granted).  On the device one sub-write per kernel launch would be all
// launch latency at NVLink rates, so several sub-writes are merged into one PIECE, released when the sub-write holding its
// LAST byte is due: the transfer never runs ahead of the reference's limiter and its last piece is released at the same
// instant as the reference's last sub-write.  Pieces end on multiples
// of 256 KiB in LAYER space (the checksum chunk), so only the fragment's own two ends touch partially covered chunks,
// exactly as for an unpaced fragment.
#pragma once
#include <cstdint>
#include <vector>

namespace dissem {
namespace host {

constexpr int64_t kPaceBurst = 256 * 1024;           // BucketSize, transport.go:408

struct PacePiece {
    int64_t off = 0;          // offset in the layer
    int64_t size = 0;
    int64_t release_ns = 0;   // earliest launch time, relative to the start of the send
};

// rate <= 0: one piece released at once.  The reference's limiter with rate 0 never grants a second sub-write
// (the send hangs); a device transfer that can never finish helps nobody, so 0 means "no limit" here.
std::vector<PacePiece> PaceSchedule(int64_t layer_off, int64_t size, int64_t rate_bytes_per_s);

}  // namespace host
}  // namespace dissem
