// Stand-in for the boost::random pieces the reference's tests use to add measurement noise
// (e.g. tests/SumKernelTest.cpp:75-77). TEST INFRASTRUCTURE ONLY (oracle build).
#ifndef GPR_ORACLE_BOOST_RANDOM_SHIM
#define GPR_ORACLE_BOOST_RANDOM_SHIM
#include <random>
namespace boost {
typedef std::minstd_rand minstd_rand;
typedef std::mt19937 mt19937;
template <class T = double>
using normal_distribution = std::normal_distribution<T>;
template <class T = double>
using uniform_real = std::uniform_real_distribution<T>;
template <class T = double>
using uniform_01 = std::uniform_real_distribution<T>;
template <class Engine, class Dist>
class variate_generator {
public:
    typedef typename Dist::result_type result_type;
    variate_generator(Engine e, Dist d) : e_(e), d_(d) {}
    result_type operator()() { return d_(e_); }
private:
    typename std::remove_reference<Engine>::type e_;
    Dist d_;
};
}  // namespace boost
#endif
