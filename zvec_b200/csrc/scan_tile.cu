// K2c placeholder until the tiled kernel lands (next commit).
#include "internal.h"
namespace zvb {
size_t scan_tile_max_parts(int sm_count) { return (size_t)sm_count * 2; }
bool scan_tile_supports(uint32_t) { return false; }
cudaError_t launch_scan_tile(const RowsView &, const QueryView &, const SearchParams &, uint64_t, uint64_t,
                             u64 *, uint32_t *, int, cudaStream_t) { return cudaErrorNotSupported; }
}  // namespace zvb
