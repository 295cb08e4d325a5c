/* TEST INFRASTRUCTURE (oracle). Not part of the product path.
 *
 * Hand-written equivalent of the header the reference's build system generates
 * at configure time (reference CMakeLists.txt:85-95 option defaults,
 * CMakeLists.txt:111-143 file template).  The values are the reference's
 * DEFAULT option states; PARALLELIZATION follows _OPENMP exactly as the
 * generated file does, DEBUGGING follows -DDEBUG.
 */
#ifndef ACCLIMATE_OPTIONS_H
#define ACCLIMATE_OPTIONS_H
#include <array>

namespace acclimate::options {

#ifdef DEBUG
constexpr bool DEBUGGING = true;
#else
constexpr bool DEBUGGING = false;
#endif
#ifdef _OPENMP
constexpr bool PARALLELIZATION = true;
#else
constexpr bool PARALLELIZATION = false;
#endif

static constexpr bool BASED_ON_INT = true;
static constexpr bool BANKERS_ROUNDING = true;
static constexpr bool CHECKPOINTING = false;
static constexpr bool CLEANUP_INFO = false;
static constexpr bool FATAL_FLOATING_POINT_EXCEPTIONS = false;
static constexpr bool FLOATING_POINT_EXCEPTIONS = false;
static constexpr bool OPTIMIZATION_PROBLEMS_FATAL = false;
static constexpr bool OPTIMIZATION_WARNINGS = false;
static constexpr bool STRICT_MIN_DERIVATIVE = false;
static constexpr bool USE_MIN_PASSAGE_IN_EXPECTATION = true;

struct Option {
    const char* name;
    bool value;
};
extern const std::array<Option, 10> options;

}  // namespace acclimate::options

#endif
