// Single-header stand-in for the subset of googletest 1.8 the reference's test
// suites use (TEST, TEST_F, TYPED_TEST_CASE / TYPED_TEST, EXPECT_* / ASSERT_*),
// so that /root/reference/test/*.cpp compile UNMODIFIED on an image without
// gtest and without network (the reference's CMake downloads gtest 1.8.0,
// CMakeLists.txt:159-165).  TEST INFRASTRUCTURE ONLY.
//
// One extension: a translation unit may define RAILS_GTEST_EXTRA_TYPES as a
// comma-led list of types before including this header; every testing::Types<...>
// list then ALSO runs those types.  That is how a new backend is added to the
// reference's `Types<StlWrapper, ...> Implementations` lists
// (test/GenericMultiVectorWrapper_test.cpp:52-58) without editing the file.
#ifndef RAILS_GTEST_SHIM_H
#define RAILS_GTEST_SHIM_H

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <sstream>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

#define GTEST_HAS_TYPED_TEST 1

#ifndef RAILS_GTEST_EXTRA_TYPES
#define RAILS_GTEST_EXTRA_TYPES
#endif

namespace testing
{

class Test
{
public:
    virtual ~Test() {}
    void Run()
    {
        SetUp();
        TestBody();
        TearDown();
    }

protected:
    virtual void SetUp() {}
    virtual void TearDown() {}
    virtual void TestBody() = 0;
};

namespace internal
{

struct Registry
{
    struct Entry
    {
        std::string name;
        Test *(*make)();
    };
    std::vector<Entry> entries;
    int failures_in_current = 0;
    static Registry &get()
    {
        static Registry r;
        return r;
    }
};

template<class T>
Test *make_test() { return new T(); }

inline bool add(std::string const &name, Test *(*make)())
{
    Registry::get().entries.push_back({name, make});
    return true;
}

template<class... Ts> struct TypeList {};

template<template<class> class Tmpl, class List> struct RegisterTyped;
template<template<class> class Tmpl>
struct RegisterTyped<Tmpl, TypeList<> >
{
    static bool run(const char *, const char *, int) { return true; }
};
template<template<class> class Tmpl, class Head, class... Tail>
struct RegisterTyped<Tmpl, TypeList<Head, Tail...> >
{
    static bool run(const char *fixture, const char *name, int index)
    {
        std::ostringstream os;
        os << fixture << "/" << index << "." << name;
        add(os.str(), &make_test<Tmpl<Head> >);
        return RegisterTyped<Tmpl, TypeList<Tail...> >::run(fixture, name, index + 1);
    }
};

// value printing when the type can be streamed, a placeholder otherwise
template<class T, class = void> struct Printer
{
    static void print(std::ostream &os, T const &) { os << "<" << sizeof(T) << "-byte object>"; }
};
template<class T>
struct Printer<T, decltype(void(std::declval<std::ostream &>() << std::declval<T const &>()))>
{
    static void print(std::ostream &os, T const &v) { os << v; }
};

inline void report(const char *file, int line, std::string const &msg)
{
    Registry::get().failures_in_current++;
    std::cout << file << ":" << line << ": Failure\n" << msg << std::endl;
}

template<class A, class B>
bool check_eq(const char *ea, const char *eb, A const &a, B const &b, const char *file, int line)
{
    if (a == b) return true;
    std::ostringstream os;
    os.precision(17);
    os << "      Expected: " << ea << " == " << eb << "\n        values: ";
    Printer<A>::print(os, a);
    os << " vs ";
    Printer<B>::print(os, b);
    report(file, line, os.str());
    return false;
}
template<class A, class B>
bool check_ne(const char *ea, const char *eb, A const &a, B const &b, const char *file, int line)
{
    if (a != b) return true;
    std::ostringstream os;
    os.precision(17);
    os << "      Expected: " << ea << " != " << eb << "\n        both are: ";
    Printer<A>::print(os, a);
    report(file, line, os.str());
    return false;
}
template<class A, class B>
bool check_gt(const char *ea, const char *eb, A const &a, B const &b, const char *file, int line)
{
    if (a > b) return true;
    std::ostringstream os;
    os.precision(17);
    os << "      Expected: " << ea << " > " << eb << "\n        values: ";
    Printer<A>::print(os, a);
    os << " vs ";
    Printer<B>::print(os, b);
    report(file, line, os.str());
    return false;
}
inline bool check_near(const char *ea, const char *eb, double a, double b, double tol, const char *file,
                       int line)
{
    if (std::fabs(a - b) <= tol) return true;
    std::ostringstream os;
    os.precision(17);
    os << "The difference between " << ea << " and " << eb << " is " << std::fabs(a - b)
       << ", which exceeds " << tol << "\n  " << ea << " = " << a << "\n  " << eb << " = " << b;
    report(file, line, os.str());
    return false;
}
// googletest's FloatingPoint<double>::AlmostEquals: at most 4 units in the last place
inline uint64_t biased(double x)
{
    uint64_t u;
    std::memcpy(&u, &x, sizeof(u));
    const uint64_t sign = (uint64_t)1 << 63;
    return (u & sign) ? ~u + 1 : (sign | u);
}
inline bool check_double_eq(const char *ea, const char *eb, double a, double b, const char *file, int line)
{
    bool ok = false;
    if (!(std::isnan(a) || std::isnan(b)))
    {
        const uint64_t x = biased(a), y = biased(b);
        ok = (x >= y ? x - y : y - x) <= 4;
    }
    if (ok) return true;
    std::ostringstream os;
    os.precision(17);
    os << "      Expected: " << ea << " and " << eb << " within 4 ULPs\n        values: " << a << " vs " << b;
    report(file, line, os.str());
    return false;
}
inline bool check_true(const char *e, bool v, bool want, const char *file, int line)
{
    if (v == want) return true;
    report(file, line, std::string("Value of: ") + e + "\n  Expected: " + (want ? "true" : "false"));
    return false;
}

}

template<class... Ts>
struct Types
{
    typedef internal::TypeList<Ts... RAILS_GTEST_EXTRA_TYPES> list;
};

inline void InitGoogleTest(int *, char **) {}

}

// Runs every registered test whose full name contains `filter` (NULL: all).
inline int RAILS_RUN_ALL_TESTS(const char *filter)
{
    using testing::internal::Registry;
    Registry &r = Registry::get();
    int ran = 0, failed = 0;
    std::vector<std::string> failed_names;
    for (auto const &e : r.entries)
    {
        if (filter && e.name.find(filter) == std::string::npos)
            continue;
        std::cout << "[ RUN      ] " << e.name << std::endl;
        r.failures_in_current = 0;
        testing::Test *t = e.make();
        t->Run();
        delete t;
        ran++;
        if (r.failures_in_current)
        {
            failed++;
            failed_names.push_back(e.name);
            std::cout << "[  FAILED  ] " << e.name << std::endl;
        }
        else
            std::cout << "[       OK ] " << e.name << std::endl;
    }
    std::cout << "[==========] " << ran << " tests ran." << std::endl;
    std::cout << "[  PASSED  ] " << ran - failed << " tests." << std::endl;
    if (failed)
    {
        std::cout << "[  FAILED  ] " << failed << " tests, listed below:" << std::endl;
        for (auto const &n : failed_names)
            std::cout << "[  FAILED  ] " << n << std::endl;
    }
    return failed ? 1 : 0;
}
#define RUN_ALL_TESTS() RAILS_RUN_ALL_TESTS(nullptr)

#define RAILS_GTEST_CAT_(a, b) a##b
#define RAILS_GTEST_CAT(a, b) RAILS_GTEST_CAT_(a, b)

#define RAILS_GTEST_PLAIN_(Suite, Name, Base)                                                   \
    class Suite##_##Name##_Test : public Base                                                   \
    {                                                                                           \
    public:                                                                                     \
        Suite##_##Name##_Test() {}                                                              \
                                                                                                \
    private:                                                                                    \
        virtual void TestBody();                                                                \
    };                                                                                          \
    static bool RAILS_GTEST_CAT(rails_gtest_reg_, __LINE__) =                                   \
        ::testing::internal::add(#Suite "." #Name,                                              \
                                 &::testing::internal::make_test<Suite##_##Name##_Test>);       \
    void Suite##_##Name##_Test::TestBody()

#define TEST(Suite, Name) RAILS_GTEST_PLAIN_(Suite, Name, ::testing::Test)
#define TEST_F(Fixture, Name) RAILS_GTEST_PLAIN_(Fixture, Name, Fixture)

#define TYPED_TEST_CASE(Fixture, TypesT) typedef TypesT::list rails_gtest_types_##Fixture

#define TYPED_TEST(Fixture, Name)                                                               \
    template<class rails_gtest_T>                                                               \
    class Fixture##_##Name##_Test : public Fixture<rails_gtest_T>                               \
    {                                                                                           \
    public:                                                                                     \
        typedef Fixture<rails_gtest_T> TestFixture;                                             \
        typedef rails_gtest_T TypeParam;                                                        \
        Fixture##_##Name##_Test() {}                                                            \
                                                                                                \
    private:                                                                                    \
        virtual void TestBody();                                                                \
    };                                                                                          \
    static bool RAILS_GTEST_CAT(rails_gtest_reg_, __LINE__) =                                   \
        ::testing::internal::RegisterTyped<Fixture##_##Name##_Test,                             \
                                           rails_gtest_types_##Fixture>::run(#Fixture, #Name, 0); \
    template<class rails_gtest_T>                                                               \
    void Fixture##_##Name##_Test<rails_gtest_T>::TestBody()

#define RAILS_GTEST_LOC_ __FILE__, __LINE__
#define EXPECT_EQ(a, b) ::testing::internal::check_eq(#a, #b, (a), (b), RAILS_GTEST_LOC_)
#define EXPECT_NE(a, b) ::testing::internal::check_ne(#a, #b, (a), (b), RAILS_GTEST_LOC_)
#define EXPECT_GT(a, b) ::testing::internal::check_gt(#a, #b, (a), (b), RAILS_GTEST_LOC_)
#define EXPECT_NEAR(a, b, tol) ::testing::internal::check_near(#a, #b, (a), (b), (tol), RAILS_GTEST_LOC_)
#define EXPECT_DOUBLE_EQ(a, b) ::testing::internal::check_double_eq(#a, #b, (a), (b), RAILS_GTEST_LOC_)
#define EXPECT_TRUE(a) ::testing::internal::check_true(#a, (bool)(a), true, RAILS_GTEST_LOC_)
#define EXPECT_FALSE(a) ::testing::internal::check_true(#a, (bool)(a), false, RAILS_GTEST_LOC_)
// fatal forms leave the test body; written `if (ok) ; else return` so that an `else`
// following the macro in user code binds to the user's own `if`
// (test/TestHelpers.hpp:27-30 has exactly that shape)
#define RAILS_GTEST_FATAL_(cond) if (cond) ; else return
#define ASSERT_EQ(a, b) RAILS_GTEST_FATAL_(EXPECT_EQ(a, b))
#define ASSERT_NE(a, b) RAILS_GTEST_FATAL_(EXPECT_NE(a, b))
#define ASSERT_GT(a, b) RAILS_GTEST_FATAL_(EXPECT_GT(a, b))
#define ASSERT_NEAR(a, b, tol) RAILS_GTEST_FATAL_(EXPECT_NEAR(a, b, tol))
#define ASSERT_DOUBLE_EQ(a, b) RAILS_GTEST_FATAL_(EXPECT_DOUBLE_EQ(a, b))
#define ASSERT_TRUE(a) RAILS_GTEST_FATAL_(EXPECT_TRUE(a))
#define ASSERT_FALSE(a) RAILS_GTEST_FATAL_(EXPECT_FALSE(a))

#endif
