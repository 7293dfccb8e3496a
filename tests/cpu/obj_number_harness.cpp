// Host harness for montecarlopathtracing_b200/csrc/obj_number.cuh (test infrastructure): one token per input line ->
// "<status_fast> <bits_fast> <status_big> <bits_big>" so that tests/test_obj_number.py can compare with glibc strtod.
#include <cstdio>
#include <cstring>
#include <iostream>
#include <string>
#include "../../montecarlopathtracing_b200/csrc/obj_number.cuh"

int main() {
    std::string tok;
    while (std::getline(std::cin, tok)) {
        double a = 0.0, b = 0.0;
        const int sa = mcpt::objnum::parse_double<false, const char*>(tok.data(), tok.data() + tok.size(), a);
        const int sb = mcpt::objnum::parse_double<true, const char*>(tok.data(), tok.data() + tok.size(), b);
        unsigned long long ua, ub;
        std::memcpy(&ua, &a, 8); std::memcpy(&ub, &b, 8);
        std::printf("%d %016llx %d %016llx %lld\n", sa, ua, sb, ub, mcpt::objnum::parse_long<const char*>(tok.data(), tok.data() + tok.size()));
    }
    return 0;
}
