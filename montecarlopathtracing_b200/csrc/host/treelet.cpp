// placeholder: replaced by the real optimiser below in a later commit
#include "host.hpp"
namespace mcpt::host {
std::vector<mcpt_bvh_node> optimize_treelet(const std::vector<mcpt_bvh_node>&) {
	throw std::runtime_error("treelet optimiser not built yet");
}
}
