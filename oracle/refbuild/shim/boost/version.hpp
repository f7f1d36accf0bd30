// Boost-over-std shim for building the reference's own hot-path sources without Boost (test infrastructure).
#pragma once
#define BOOST_VERSION 107700
