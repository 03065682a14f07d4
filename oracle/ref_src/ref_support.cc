// TEST INFRASTRUCTURE -- what the reference's enforce.cc expects from the rest of libcommon when
// it is compiled on its own for oracle/_ref: the one exported flag it reads
// (PHI_DEFINE_EXPORTED_int32(call_stack_level, ...) in paddle/common/flags.cc).
#include <stdint.h>

namespace paddle_flags {
__attribute__((visibility("default"))) int32_t FLAGS_call_stack_level = 1;
}  // namespace paddle_flags
