// Debug / A-B switches.  Release builds (the default, what ships in lib/libq3tts_decode.so) read NO environment variables and
// compile every `dbg` branch out of the kernels; `python -m qwen3_tts_b200.build --debug` builds lib/libq3tts_decode_dbg.so with
// -DQ3TTS_DEBUG, where the Q3TTS_* switches and the GemmParams::dbg experiment bits used by tools/ exist.
#pragma once
#include <cstdlib>

#ifdef Q3TTS_DEBUG
#define Q3_ENV(name) getenv(name)
#define Q3_DBG(p, bits) (((p).dbg & (bits)) != 0)
#else
#define Q3_ENV(name) (static_cast<const char*>(nullptr))
#define Q3_DBG(p, bits) (false)
#endif
