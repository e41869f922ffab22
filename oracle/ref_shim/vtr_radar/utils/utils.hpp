// TEST INFRASTRUCTURE ONLY.  Stand-in for vtr_radar/utils/utils.hpp: detector.{hpp,inl} include it
// but name nothing from it.
#pragma once
