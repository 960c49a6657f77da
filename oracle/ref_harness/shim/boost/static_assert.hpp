// Stand-in for <boost/static_assert.hpp> (oracle build only; boost is not installed).
#pragma once
#define BOOST_STATIC_ASSERT(x) static_assert(x, #x)
