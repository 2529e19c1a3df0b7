// srt_host_context.h -- the "current context" of the host wrappers.
//
// The reference's loaders talk to whatever GL context / compute program is current on the calling thread
// ("compute shader program should be bound before calling", include/asset_utils/gpu_loader.h:17,24).  The
// wrappers keep that calling convention over the C-ABI of include/srt_b200.h: Graphics::Compute::Use() makes its
// srt_ctx current, and AssetUtils::Upload*/Graphics::Texture* act on the current one.
#pragma once

#include <stdexcept>
#include <string>

#include "srt_b200.h"

namespace SrtHost {

srt_ctx* Current();             // throws std::runtime_error when no Compute program is in use
srt_ctx* CurrentOrNull();
void MakeCurrent(srt_ctx* ctx);

// turns a non-zero C-ABI status into the exception the reference-shaped wrappers throw
inline void Check(int status, srt_ctx* ctx, const char* what) {
  if (status != 0) throw std::runtime_error(std::string(what) + ": " + srt_last_error(ctx));
}

}  // namespace SrtHost
