// TEST INFRASTRUCTURE: stand-in; the reference only uses this_thread::sleep inside its viewers (grid.cpp:309).
#ifndef REF_STANDIN_BOOST_THREAD
#define REF_STANDIN_BOOST_THREAD
#include "boost/shared_ptr.hpp"
namespace boost {
namespace posix_time { struct microseconds { long us; explicit microseconds(long u) : us(u) {} }; }
namespace this_thread { inline void sleep(const posix_time::microseconds&) {} }
}  // namespace boost
#endif
