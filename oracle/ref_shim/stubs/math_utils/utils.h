// ORACLE — TEST INFRASTRUCTURE ONLY.  Stand-in for libs/math_utils/utils.h, whose matrix headers are MSVC-only
// (template parameter shadowing, rejected by gcc).  Declares the two functions conegaussiansampler.cpp uses; their
// bodies are the reference's own lines, extracted at build time from libs/math_utils/utils.cpp:149-165 into
// oracle/_ref/rodrigues_excerpt.inc (see oracle/Makefile) and compiled inside ref_api.cpp.
#pragma once
#include <glm/glm.hpp>
glm::vec3 RodriguesRotation(glm::vec3 v, float teta, glm::vec3 k);
glm::dvec3 RodriguesRotation(glm::dvec3 v, double teta, glm::dvec3 k);
// DEGREE_TO_RADIANS: the reference's own macro, extracted at build time (oracle/Makefile)
#include <glm/gtc/constants.hpp>
#include <degree_to_radians_excerpt.inc>
