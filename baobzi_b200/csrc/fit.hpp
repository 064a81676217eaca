// fit.hpp -- host fitter behind baobzi_init.  The user's callback runs on the host, so fitting stays
// there (BASELINE.json north_star); its product is a HostTree that is then flattened to the GPU.
#pragma once

#include "../../include/baobzi.h"
#include "host_tree.hpp"

#include <string>

namespace bz {

// Adaptive piecewise-Chebyshev fit of input.func on the box center +- half_length, following the
// reference's algorithm (include/baobzi_template.hpp:226-360 Node::fit, 406-457 FunctionTree,
// 650-760 Function).  Returns false with a message when max_depth is exceeded.
bool fit_tree(const baobzi_input_t &input, const double *center, const double *half_length, HostTree &out,
              std::string &err);

} // namespace bz
