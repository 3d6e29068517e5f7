// Test infrastructure only (oracle/_ref build). Link-time stub for three
// ailego::BufferHandle members declared at
// src/include/zvec/ailego/buffer/buffer_manager.h:203-246; the real ones live in
// buffer_manager.cc, which needs Arrow/Parquet and is not built. They are never
// executed with the mmap / memory storages the oracle uses.
#include <cstdlib>
namespace arrow { class ChunkedArray; class Buffer; }
#include <zvec/ailego/buffer/buffer_manager.h>
namespace zvec { namespace ailego {
BufferHandle::~BufferHandle() {}
void *BufferHandle::pin_vector_data() { std::abort(); }
bool BufferHandle::unpin_vector_data() { std::abort(); }
} }
