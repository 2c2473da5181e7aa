// Stand-in for <mrpt/core/optional_ref.h>
#pragma once
#include <functional>
#include <optional>
namespace mrpt {
template <class T> using optional_ref = std::optional<std::reference_wrapper<T>>;
}
