// Shared between capi_host.cpp (host-only entry points) and engine.cu (numeric entry points).
#pragma once
#include <string>

#include "symbolic.h"

struct spchol_symbolic { spchol::Symbolic S; };

namespace spchol {
void set_error(const std::string& msg);
inline const Symbolic& symbolic_of(const spchol_symbolic* s) { return s->S; }
}  // namespace spchol
