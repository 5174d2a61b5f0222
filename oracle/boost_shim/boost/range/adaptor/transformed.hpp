// Stub for <boost/range/adaptor/transformed.hpp>: the reference only has a bare
// `using boost::adaptors::transformed;` (src/pairwise.cpp:13) and never uses it.
#pragma once
namespace boost { namespace adaptors { struct transformed_stub_t {}; static const transformed_stub_t transformed{}; } }
