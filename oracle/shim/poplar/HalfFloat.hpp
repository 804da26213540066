// TEST INFRASTRUCTURE ONLY (oracle/). Empty stand-in for <poplar/HalfFloat.hpp>
// (included by /root/reference/src/codelets/codelets.cpp:7); `half` comes from
// the Vertex.hpp shim (cuda_fp16.h host implementation, round-to-nearest-even).
#pragma once
