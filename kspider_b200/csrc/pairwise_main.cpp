// `pairwise <index_prefix> <threads>` — same command line as the reference binary (pairwise.cpp:3-5).
#include <cstdio>
#include <exception>
#include <string>

#include "kspider_dropin.hpp"

int main(int argc, char** argv) {
    if (argc < 3) { std::fprintf(stderr, "usage: %s <index_prefix> <threads>\n", argv[0]); return 2; }
    try {
        kSpider::pairwise(argv[1], std::stoi(argv[2]));
    } catch (const std::exception& e) {
        std::fprintf(stderr, "%s\n", e.what());
        return 1;
    }
    return 0;
}
