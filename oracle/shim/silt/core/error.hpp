#pragma once
#include <silt/core/types.hpp>
namespace silt { namespace error {
struct mismatch_host : std::invalid_argument {
  mismatch_host(host_t, host_t) : std::invalid_argument("tensor host mismatch") {}
};
struct missing_file : std::invalid_argument {
  missing_file(const std::string& s) : std::invalid_argument("missing file " + s) {}
};
}}  // namespace silt::error
