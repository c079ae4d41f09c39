// Stand-in for the reference's missing progress-bar submodule (C++/lib/indicators, status "missing"
// in .SUBMODULES.json), included by C++/src/viewport.cpp:2. Only the names that file uses; every
// call is a no-op. Test infrastructure for oracle/_ref (see ref_cpp_shim.cpp).
#pragma once
#include <cstddef>
#include <vector>
namespace indicators {
enum class Color { green };
enum class FontStyle { bold };
namespace option {
struct BarWidth { BarWidth(int) {} };
struct Start { Start(const char*) {} };
struct Fill { Fill(const char*) {} };
struct Lead { Lead(const char*) {} };
struct Remainder { Remainder(const char*) {} };
struct End { End(const char*) {} };
struct PrefixText { PrefixText(const char*) {} };
struct ForegroundColor { ForegroundColor(Color) {} };
struct ShowElapsedTime { ShowElapsedTime(bool) {} };
struct ShowRemainingTime { ShowRemainingTime(bool) {} };
struct FontStyles { FontStyles(std::vector<FontStyle>) {} };
struct MaxProgress { MaxProgress(std::size_t) {} };
}  // namespace option
struct ProgressBar {
    template <class... A> ProgressBar(A&&...) {}
    void set_progress(std::size_t) {}
    void mark_as_completed() {}
};
}  // namespace indicators
