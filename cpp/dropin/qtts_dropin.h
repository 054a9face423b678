// qtts_dropin.h — force-included (g++ -include) in front of the reference's OWN translation units
// (src/qwen3_tts.cpp, src/main.cpp, compiled where they lie) to build the reference pipeline and CLI
// on top of qtts_b200::TTSTransformer without editing a line of the reference.
//
// It is the INTEGRATION.md section 1 patch (`using TTSTransformerImpl = qtts_b200::TTSTransformer;`
// + the member's type) done from the outside:
//   1. the reference's tts_transformer.h is included first, under its original names (its
//      `#pragma once` then makes the include in src/qwen3_tts.h:4 a no-op);
//   2. the adapter is included;
//   3. from here on the token `TTSTransformer` names an alias of qtts_b200::TTSTransformer inside
//      namespace qwen3_tts, so `TTSTransformer transformer_;` (src/qwen3_tts.h:132) and
//      `TTSTransformer::frame_callback_t` (src/qwen3_tts.cpp:438,455) bind to the B200 engine.
// The reference's own class stays declared (unused); src/tts_transformer.cpp is NOT compiled.
#pragma once
#include "tts_transformer.h"            // the reference's (found through -I<reference>/src)
#include "qtts_transformer.hpp"
namespace qwen3_tts { using TTSTransformerB200 = qtts_b200::TTSTransformer; }
#define TTSTransformer TTSTransformerB200
