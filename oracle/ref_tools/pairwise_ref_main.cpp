// Entry point for oracle/_ref/pairwise_ref: calls the reference's own kSpider::pairwise
// (compiled from /root/reference/src/pairwise.cpp where it lies) exactly like the reference's
// `pairwise` binary does (pairwise.cpp:3-5).  kSpider.hpp is not included because it drags in
// kProcessor; the one declaration needed is repeated here.
#include <string>
namespace kSpider { void pairwise(std::string index_prefix, int user_threads); }
int main(int argc, char** argv) {
    if (argc < 3) return 2;
    kSpider::pairwise(argv[1], std::stoi(argv[2]));
    return 0;
}
