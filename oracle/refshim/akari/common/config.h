// Hand-written stand-in for the header the reference GENERATES at build time
// (tools/configure.cpp:44-103 -> src/akari/common/config.h). Only the "rgb" variant
// (resources/akari.conf:20-23) is enabled, exactly as in the reference's default build.
// TEST INFRASTRUCTURE ONLY (oracle/_ref build); not part of the product path.
#pragma once
#include <type_traits>
#include <string_view>
#include <stdexcept>
namespace akari {
#define AKR_CORE_STRUCT(Name) template struct Name<float>;
#define AKR_CORE_CLASS(Name) template class Name<float>;
#define AKR_RENDER_STRUCT(Name) template struct Name<Config<float, Color<float, 3>>>;
#define AKR_RENDER_CLASS(Name) template class Name<Config<float, Color<float, 3>>>;
template <typename C> constexpr const char *get_variant_string() { return "rgb"; }
#define AKR_INVOKE_VARIANT(variant, func, ...) ([&]() { return func<Config<float, Color<float, 3>>>(__VA_ARGS__); })()
constexpr const char *default_variant = "rgb";
constexpr const char *enabled_variants[] = {"rgb"};
} // namespace akari
