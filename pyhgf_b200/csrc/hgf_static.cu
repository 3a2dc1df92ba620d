#include "hgf_dispatch.h"
namespace hgf {
static const KernelVariant* const kNone[] = {nullptr};
const KernelVariant* const* static_variants() { return kNone; }
}
