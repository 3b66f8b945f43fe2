// ngs_parse.cpp — host-side parsing of the two strings the metric sync hands to the dealer
// (/root/reference/pkg/controller/node.go:111-135): the usage value formatted with %.5f
// (pkg/prometheus/prometheus.go:61) and the update time "2006-01-02T15:04:05Z" rendered in
// Asia/Shanghai (pkg/controller/node.go:137-141, pkg/dealer/type.go:12, stats.go:38-44).
#include <cerrno>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>

#include "../../include/ngs_dealer.h"

namespace {
bool two_digits(const char* s, int* v) {
    if (s[0] < '0' || s[0] > '9' || s[1] < '0' || s[1] > '9') return false;
    *v = (s[0] - '0') * 10 + (s[1] - '0');
    return true;
}
// days since 1970-01-01 of a proleptic Gregorian date
int64_t days_from_civil(int64_t y, int m, int d) {
    y -= m <= 2;
    const int64_t era = (y >= 0 ? y : y - 399) / 400;
    const int64_t yoe = y - era * 400;
    const int64_t doy = (153 * (m + (m > 2 ? -3 : 9)) + 2) / 5 + d - 1;
    const int64_t doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
    return era * 146097 + doe - 719468;
}
}  // namespace

extern "C" {

// strconv.ParseFloat(s, 64) as GetUsage uses it (nodeusage.go:103): the whole string must be a
// number, no surrounding blanks; a range error counts as failure (err != nil).
int ngs_parse_usage(const char* s, double* value_out) {
    if (!s || !*s || !value_out) return 0;
    const unsigned char c0 = (unsigned char)s[0];
    if (c0 == ' ' || c0 == '\t' || c0 == '\n' || c0 == '\v' || c0 == '\f' || c0 == '\r') return 0;
    errno = 0;
    char* end = nullptr;
    const double v = strtod(s, &end);
    if (end == s || *end != '\0') return 0;
    if (errno == ERANGE && std::isinf(v)) return 0;
    *value_out = v;
    return 1;
}

// time.ParseInLocation("2006-01-02T15:04:05Z", s, Asia/Shanghai) (stats.go:32-44).  len < 5 fails
// first (stats.go:32-35).  The trailing 'Z' of that layout is a literal, the zone comes from the
// location: UTC+8 (no DST since 1991; earlier timestamps are outside what the sync ever writes).
// A fractional second after the seconds field is accepted, as Go does, and dropped.
int ngs_parse_update_time(const char* s, int64_t* ts_out) {
    if (!s || !ts_out) return 0;
    const size_t len = strlen(s);
    if (len < 5) return 0;
    if (len < 20) return 0;
    int Y1, Y2, mo, d, hh, mi, ss;
    if (!two_digits(s, &Y1) || !two_digits(s + 2, &Y2) || s[4] != '-' || !two_digits(s + 5, &mo) || s[7] != '-' ||
        !two_digits(s + 8, &d) || s[10] != 'T' || !two_digits(s + 11, &hh) || s[13] != ':' || !two_digits(s + 14, &mi) ||
        s[16] != ':' || !two_digits(s + 17, &ss))
        return 0;
    size_t i = 19;
    if (s[i] == '.' || s[i] == ',') {  // optional fractional seconds
        size_t j = i + 1;
        while (s[j] >= '0' && s[j] <= '9') j++;
        if (j == i + 1) return 0;
        i = j;
    }
    if (s[i] != 'Z' || s[i + 1] != '\0') return 0;
    const int year = Y1 * 100 + Y2;
    if (mo < 1 || mo > 12 || hh > 23 || mi > 59 || ss > 59) return 0;
    static const int mdays[12] = {31, 28, 31, 30, 31, 30, 31, 31, 30, 31, 30, 31};
    const bool leap = (year % 4 == 0 && year % 100 != 0) || year % 400 == 0;
    const int dim = mdays[mo - 1] + (mo == 2 && leap ? 1 : 0);
    if (d < 1 || d > dim) return 0;
    *ts_out = days_from_civil(year, mo, d) * 86400 + hh * 3600 + mi * 60 + ss - 8 * 3600;
    return 1;
}

}  // extern "C"
