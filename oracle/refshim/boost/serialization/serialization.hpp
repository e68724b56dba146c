// Stand-in for the two boost::serialization names DBoW2's BowVector.h / FeatureVector.h mention (TEST INFRASTRUCTURE).
// Their `serialize` member templates are never instantiated by the transform path.
#pragma once
namespace boost { namespace serialization {
class access {};
template <class Base, class Derived> Base& base_object(Derived& d) { return d; }
}}
