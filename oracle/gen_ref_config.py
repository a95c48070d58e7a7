"""Writes the generated header the reference's build would produce with resources/configure.py
for a single `scalar_rgb` variant (reference: resources/configure.py:17-110).  Output goes to
oracle/_ref/ only (git-ignored build artefact)."""
import os, sys
out = sys.argv[1]
os.makedirs(os.path.dirname(out), exist_ok=True)
T = "float, Color<float, 3>"
open(out, "w").write(f"""#pragma once
#include <mitsuba/core/fwd.h>
#define MI_VARIANTS "scalar_rgb\\n"
#define MI_DEFAULT_VARIANT "scalar_rgb"
#define MI_EXTERN_STRUCT(Name) extern template struct Name<{T}>;
#define MI_EXTERN_CLASS(Name) extern template class Name<{T}>;
#define MI_INSTANTIATE_STRUCT(Name) template struct Name<{T}>;
#define MI_INSTANTIATE_CLASS(Name) template class Name<{T}>;
#define MI_INVOKE_VARIANT(variant, func, ...) [&]() {{ return func<{T}>(__VA_ARGS__); }}()
NAMESPACE_BEGIN(mitsuba)
NAMESPACE_BEGIN(detail)
template <typename Float_, typename Spectrum_> constexpr const char *get_variant() {{ return "scalar_rgb"; }}
NAMESPACE_END(detail)
NAMESPACE_END(mitsuba)
""")
