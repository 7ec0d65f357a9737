// Small restatements of the Go standard-library behaviours the reference's observable output depends on:
// strings.TrimSpace (Unicode), strconv.ParseBool, time.ParseDuration / Duration.String, log's line prefix.
#pragma once
#include <cstdint>
#include <string>

namespace gocompat {

// strings.TrimSpace over a Go string holding arbitrary bytes (used by internal/lm/machine-type.go:50 and
// internal/vgpu/pciutil.go:86).  Invalid UTF-8 decodes to RuneError (width 1), which is not a space.
std::string trim_space(const std::string &s);

// strconv.ParseBool: 1 t T TRUE true True / 0 f F FALSE false False
bool parse_bool(const std::string &s, bool *out);

// time.ParseDuration; result in nanoseconds.  Returns false on a syntax error / overflow.
bool parse_duration(const std::string &s, int64_t *ns);
// time.Duration.String(): "1m0s", "1.5s", "500ms", "0s", ...
std::string duration_string(int64_t ns);

// strings.Replace(s, " ", "-", -1)
std::string spaces_to_dashes(const std::string &s);

// log.Printf with the default flags: "2006/01/02 15:04:05 " prefix on stderr, newline appended if missing.
void logf(const char *fmt, ...) __attribute__((format(printf, 1, 2)));

}  // namespace gocompat
