#pragma once
#include <mitsuba/core/fwd.h>
#define MI_VARIANTS "scalar_rgb\n"
#define MI_DEFAULT_VARIANT "scalar_rgb"
#define MI_EXTERN_STRUCT(Name) extern template struct Name<float, Color<float, 3>>;
#define MI_EXTERN_CLASS(Name) extern template class Name<float, Color<float, 3>>;
#define MI_INSTANTIATE_STRUCT(Name) template struct Name<float, Color<float, 3>>;
#define MI_INSTANTIATE_CLASS(Name) template class Name<float, Color<float, 3>>;
#define MI_INVOKE_VARIANT(variant, func, ...) [&]() { return func<float, Color<float, 3>>(__VA_ARGS__); }()
NAMESPACE_BEGIN(mitsuba)
NAMESPACE_BEGIN(detail)
template <typename Float_, typename Spectrum_> constexpr const char *get_variant() { return "scalar_rgb"; }
NAMESPACE_END(detail)
NAMESPACE_END(mitsuba)
