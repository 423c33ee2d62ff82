// TEST INFRASTRUCTURE ONLY -- minimal stand-in for <boost/serialization/*.hpp> so that the reference headers under
// /root/reference compile without Boost (oracle/Makefile.ref). The reference only befriends boost::serialization::access
// and names base_object / BOOST_SERIALIZATION_NVP inside never-instantiated serialize() member templates.
#pragma once
#include <iostream>
namespace boost { namespace serialization {
class access;
template <class Base, class Derived> Base& base_object(Derived& d) { return static_cast<Base&>(d); }
template <class T> T& make_nvp(const char*, T& t) { return t; }
}}
#ifndef BOOST_SERIALIZATION_NVP
#define BOOST_SERIALIZATION_NVP(x) x
#endif
