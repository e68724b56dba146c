// TEST INFRASTRUCTURE.  Force-included FIRST by the builds around refshim/ref_objects_capi.cc: the wrappers set state that the
// reference's Frame / KeyFrame / MapPoint keep private or protected (AssignFeaturesToGrid, mDescriptor, mbBad, mObservations ...).
// The standard headers are included before the keywords are redefined, so only ORB-SLAM3's (and the stand-ins') class bodies are
// affected: access changes, layouts do not.
#pragma once
#include <algorithm>
#include <atomic>
#include <cassert>
#include <chrono>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <iostream>
#include <list>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <sstream>
#include <string>
#include <thread>
#include <tuple>
#include <unordered_map>
#include <unordered_set>
#include <utility>
#include <vector>
#define private public
#define protected public
