// Stub for <boost/algorithm/string/join.hpp>: the reference only has a bare
// `using boost::algorithm::join;` (src/pairwise.cpp:14) and never calls it.
#pragma once
namespace boost { namespace algorithm { struct join_stub_t {}; static const join_stub_t join{}; } }
