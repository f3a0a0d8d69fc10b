// TEST INFRASTRUCTURE: stands in for Thirdparty/DBoW2/include/DBoW2/TemplatedVocabulary.h (which needs
// cv::FileStorage etc.).  The reference's front-end only names the type (Features/ORBVocabulary.hpp) and
// passes a null pointer (demo/demo_initialization.cpp:76-77).  BowVector / FeatureVector are the real headers.
#pragma once
#include "BowVector.h"
#include "FeatureVector.h"
namespace DBoW2 {
template <class TDescriptor, class F>
class TemplatedVocabulary {};
}  // namespace DBoW2
