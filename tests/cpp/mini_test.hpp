// mini_test.hpp -- a small stand-in for the slice of GoogleTest the reference's operator tests
// use (TEST_F, EXPECT_*): gtest is not installed in this image and there is no network.
#pragma once
#include <cmath>
#include <functional>
#include <iostream>
#include <string>
#include <vector>

namespace mini_test {
struct Case {
    std::string name;
    std::function<void(int&)> run;
};
inline std::vector<Case>& registry() {
    static std::vector<Case> r;
    return r;
}
struct Registrar {
    Registrar(const std::string& name, std::function<void(int&)> fn) { registry().push_back({name, std::move(fn)}); }
};
inline int run_all() {
    int failed_cases = 0;
    for (auto& c : registry()) {
        int failures = 0;
        std::cout << "[ RUN      ] " << c.name << std::endl;
        try {
            c.run(failures);
        } catch (const std::exception& e) {
            ++failures;
            std::cout << "  exception: " << e.what() << std::endl;
        }
        std::cout << (failures ? "[  FAILED  ] " : "[       OK ] ") << c.name << std::endl;
        failed_cases += failures ? 1 : 0;
    }
    std::cout << registry().size() - failed_cases << " of " << registry().size() << " tests passed" << std::endl;
    return failed_cases ? 1 : 0;
}
}  // namespace mini_test

#define MINI_CHECK(cond, text)                                                              \
    do {                                                                                    \
        if (!(cond)) {                                                                      \
            ++mini_failures__;                                                              \
            std::cout << "  " << __FILE__ << ":" << __LINE__ << ": expected " << text << std::endl; \
        }                                                                                   \
    } while (0)
#define TEST_F(fixture, name)                                                               \
    struct fixture##_##name : fixture {                                                     \
        void body(int& mini_failures__);                                                    \
    };                                                                                      \
    static mini_test::Registrar reg_##fixture##_##name(#fixture "." #name, [](int& f) {     \
        fixture##_##name t;                                                                 \
        t.SetUp(f);                                                                         \
        if (!f) t.body(f);                                                                  \
    });                                                                                     \
    void fixture##_##name::body(int& mini_failures__)
#define EXPECT_TRUE(c) MINI_CHECK((c), #c)
#define EXPECT_FALSE(c) MINI_CHECK(!(c), "!(" #c ")")
#define EXPECT_GT(a, b) MINI_CHECK((a) > (b), #a " > " #b)
#define EXPECT_LT(a, b) MINI_CHECK((a) < (b), #a " < " #b)
#define EXPECT_NEAR(a, b, tol) MINI_CHECK(std::fabs((a) - (b)) <= (tol), #a " ~= " #b)
#define MINI_TEST_MAIN() \
    int main() { return mini_test::run_all(); }
