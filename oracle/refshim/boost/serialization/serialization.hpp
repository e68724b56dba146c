// Stand-in for the two boost::serialization names DBoW2's BowVector.h / FeatureVector.h mention (TEST INFRASTRUCTURE).
// Their `serialize` member templates are never instantiated by the transform path.
#pragma once
namespace boost { namespace serialization {
class access {};
template <class Base, class Derived> Base& base_object(Derived& d) { return d; }
}}
// camera models: BOOST_CLASS_EXPORT_IMPLEMENT(...) / BOOST_SERIALIZATION_ASSUME_ABSTRACT(...) at namespace scope
#ifndef BOOST_CLASS_EXPORT_IMPLEMENT
#define BOOST_CLASS_EXPORT_IMPLEMENT(T)
#define BOOST_CLASS_EXPORT_KEY(T)
#define BOOST_CLASS_EXPORT(T)
#define BOOST_SERIALIZATION_ASSUME_ABSTRACT(T)
#define BOOST_CLASS_VERSION(T, N)
#endif
