#!/usr/bin/env python
"""Writes tests/golden/upstream_robfig.json: known-answer rows of robfig/cron v3.0.1's OWN tests (spec_test.go
TestNext / TestActivation, parser_test.go TestRange / TestParseScheduleErrors / TestStandardSpecSchedule).

[UPSTREAM-RECALLED]  The module is pinned by the reference (go.mod:10, imported at
internal/controller/cron_controller.go:28) but its source is not under /root/reference and no Go toolchain exists
here, so these rows were written down from knowledge of the public upstream test files -- they are NOT the output of a
Go run, and NOT generated from this repo's oracle (this script never imports it).  Their value: they are the only
vectors for zones / DST gaps and overlaps / the DOM-DOW rule / Feb 29 / unsatisfiable specs whose expected values come
from outside this repository.

Rows whose upstream time string carries no zone ("Mon Jul 9 14:45 2012") are parsed by upstream in time.Local and
compared in time.Local; they are evaluated in UTC here (no DST, so the wall-clock arithmetic is identical).  Rows
"TZ=America/New_York 2012-11-04T..." put the *time* in that location with a spec that has no TZ= prefix: upstream then
evaluates in the time's own location (spec.go: `if loc == time.Local { loc = t.Location() }`), i.e. `zone` below.
TestNext's parser is NewParser(Second|Minute|Hour|Dom|Month|DowOptional|Descriptor): every row has 6 fields, which
b200cron_parse_with_seconds accepts.  TestActivation uses ParseStandard (5 fields) and checks
`Next(t - 1s) == t` against `expected`.
"""
import datetime as dt
import json
import os

UTC = dt.timezone.utc


def epoch(s: str) -> int:
    if s[:4].isdigit():  # 2012-03-11T00:00:00-0500
        return int(dt.datetime.strptime(s, "%Y-%m-%dT%H:%M:%S%z").timestamp())
    parts = s.split()  # "Mon Jul 9 14:45 2012" / "Mon Jul 9 14:59:59 2012" (Go ignores the weekday's consistency)
    hms = parts[3] if parts[3].count(":") == 2 else parts[3] + ":00"
    d = dt.datetime.strptime(f"{parts[1]} {parts[2]} {hms} {parts[4]}", "%b %d %H:%M:%S %Y")
    return int(d.replace(tzinfo=UTC).timestamp())


NY, SP = "America/New_York", "America/Sao_Paulo"
J = "Mon Jul 9 "
NEXT = [  # (time, spec, expected, zone the schedule is evaluated in)
    # Simple cases
    (J + "14:45 2012", "0 0/15 * * * *", J + "15:00 2012", "UTC"),
    (J + "14:59 2012", "0 0/15 * * * *", J + "15:00 2012", "UTC"),
    (J + "14:59:59 2012", "0 0/15 * * * *", J + "15:00 2012", "UTC"),
    # Wrap around hours
    (J + "15:45 2012", "0 20-35/15 * * * *", J + "16:20 2012", "UTC"),
    # Wrap around days
    (J + "23:46 2012", "0 */15 * * * *", "Tue Jul 10 00:00 2012", "UTC"),
    (J + "23:45 2012", "0 20-35/15 * * * *", "Tue Jul 10 00:20 2012", "UTC"),
    (J + "23:35:51 2012", "15/35 20-35/15 * * * *", "Tue Jul 10 00:20:15 2012", "UTC"),
    (J + "23:35:51 2012", "15/35 20-35/15 1/2 * * *", "Tue Jul 10 01:20:15 2012", "UTC"),
    (J + "23:35:51 2012", "15/35 20-35/15 10-12 * * *", "Tue Jul 10 10:20:15 2012", "UTC"),
    (J + "23:35:51 2012", "15/35 20-35/15 1/2 */2 * *", "Thu Jul 11 01:20:15 2012", "UTC"),
    (J + "23:35:51 2012", "15/35 20-35/15 * 9-20 * *", "Wed Jul 10 00:20:15 2012", "UTC"),
    (J + "23:35:51 2012", "15/35 20-35/15 * 9-20 Jul *", "Wed Jul 10 00:20:15 2012", "UTC"),
    # Wrap around months
    (J + "23:35 2012", "0 0 0 9 Apr-Oct ?", "Thu Aug 9 00:00 2012", "UTC"),
    (J + "23:35 2012", "0 0 0 */5 Apr,Aug,Oct Mon", "Tue Aug 1 00:00 2012", "UTC"),   # DOM/DOW OR rule
    (J + "23:35 2012", "0 0 0 */5 Oct Mon", "Mon Oct 1 00:00 2012", "UTC"),
    # Wrap around years
    (J + "23:35 2012", "0 0 0 * Feb Mon", "Mon Feb 4 00:00 2013", "UTC"),              # AND rule (dom is `*`)
    (J + "23:35 2012", "0 0 0 * Feb Mon/2", "Fri Feb 1 00:00 2013", "UTC"),
    # Wrap around minute, hour, day, month, and year
    ("Mon Dec 31 23:59:45 2012", "0 * * * * *", "Tue Jan 1 00:00:00 2013", "UTC"),
    # Leap year
    (J + "23:35 2012", "0 0 0 29 Feb ?", "Mon Feb 29 00:00 2016", "UTC"),
    # Daylight savings time 2am EST (-5) -> 3am EDT (-4)
    ("2012-03-11T00:00:00-0500", "TZ=America/New_York 0 30 2 11 Mar ?", "2013-03-11T02:30:00-0400", NY),
    # hourly job
    ("2012-03-11T00:00:00-0500", "TZ=America/New_York 0 0 * * * ?", "2012-03-11T01:00:00-0500", NY),
    ("2012-03-11T01:00:00-0500", "TZ=America/New_York 0 0 * * * ?", "2012-03-11T03:00:00-0400", NY),
    ("2012-03-11T03:00:00-0400", "TZ=America/New_York 0 0 * * * ?", "2012-03-11T04:00:00-0400", NY),
    ("2012-03-11T04:00:00-0400", "TZ=America/New_York 0 0 * * * ?", "2012-03-11T05:00:00-0400", NY),
    # hourly job using CRON_TZ
    ("2012-03-11T00:00:00-0500", "CRON_TZ=America/New_York 0 0 * * * ?", "2012-03-11T01:00:00-0500", NY),
    ("2012-03-11T01:00:00-0500", "CRON_TZ=America/New_York 0 0 * * * ?", "2012-03-11T03:00:00-0400", NY),
    ("2012-03-11T03:00:00-0400", "CRON_TZ=America/New_York 0 0 * * * ?", "2012-03-11T04:00:00-0400", NY),
    ("2012-03-11T04:00:00-0400", "CRON_TZ=America/New_York 0 0 * * * ?", "2012-03-11T05:00:00-0400", NY),
    # 1am nightly job
    ("2012-03-11T00:00:00-0500", "TZ=America/New_York 0 0 1 * * ?", "2012-03-11T01:00:00-0500", NY),
    ("2012-03-11T01:00:00-0500", "TZ=America/New_York 0 0 1 * * ?", "2012-03-12T01:00:00-0400", NY),
    # 2am nightly job (skipped)
    ("2012-03-11T00:00:00-0500", "TZ=America/New_York 0 0 2 * * ?", "2012-03-12T02:00:00-0400", NY),
    # Daylight savings time 2am EDT (-4) => 1am EST (-5)
    ("2012-11-04T00:00:00-0400", "TZ=America/New_York 0 30 2 04 Nov ?", "2012-11-04T02:30:00-0500", NY),
    ("2012-11-04T01:45:00-0400", "TZ=America/New_York 0 30 1 04 Nov ?", "2012-11-04T01:30:00-0500", NY),
    # hourly job
    ("2012-11-04T00:00:00-0400", "TZ=America/New_York 0 0 * * * ?", "2012-11-04T01:00:00-0400", NY),
    ("2012-11-04T01:00:00-0400", "TZ=America/New_York 0 0 * * * ?", "2012-11-04T01:00:00-0500", NY),
    ("2012-11-04T01:00:00-0500", "TZ=America/New_York 0 0 * * * ?", "2012-11-04T02:00:00-0500", NY),
    # 1am nightly job (runs twice)
    ("2012-11-04T00:00:00-0400", "TZ=America/New_York 0 0 1 * * ?", "2012-11-04T01:00:00-0400", NY),
    ("2012-11-04T01:00:00-0400", "TZ=America/New_York 0 0 1 * * ?", "2012-11-04T01:00:00-0500", NY),
    ("2012-11-04T01:00:00-0500", "TZ=America/New_York 0 0 1 * * ?", "2012-11-05T01:00:00-0500", NY),
    # 2am nightly job
    ("2012-11-04T00:00:00-0400", "TZ=America/New_York 0 0 2 * * ?", "2012-11-04T02:00:00-0500", NY),
    ("2012-11-04T02:00:00-0500", "TZ=America/New_York 0 0 2 * * ?", "2012-11-05T02:00:00-0500", NY),
    # 3am nightly job
    ("2012-11-04T00:00:00-0400", "TZ=America/New_York 0 0 3 * * ?", "2012-11-04T03:00:00-0500", NY),
    ("2012-11-04T03:00:00-0500", "TZ=America/New_York 0 0 3 * * ?", "2012-11-05T03:00:00-0500", NY),
    # the same with the time (not the spec) carrying the location
    ("2012-11-04T00:00:00-0400", "0 0 * * * ?", "2012-11-04T01:00:00-0400", NY),
    ("2012-11-04T01:00:00-0400", "0 0 * * * ?", "2012-11-04T01:00:00-0500", NY),
    ("2012-11-04T01:00:00-0500", "0 0 * * * ?", "2012-11-04T02:00:00-0500", NY),
    ("2012-11-04T00:00:00-0400", "0 0 1 * * ?", "2012-11-04T01:00:00-0400", NY),
    ("2012-11-04T01:00:00-0400", "0 0 1 * * ?", "2012-11-04T01:00:00-0500", NY),
    ("2012-11-04T01:00:00-0500", "0 0 1 * * ?", "2012-11-05T01:00:00-0500", NY),
    ("2012-11-04T00:00:00-0400", "0 0 2 * * ?", "2012-11-04T02:00:00-0500", NY),
    ("2012-11-04T02:00:00-0500", "0 0 2 * * ?", "2012-11-05T02:00:00-0500", NY),
    ("2012-11-04T00:00:00-0400", "0 0 3 * * ?", "2012-11-04T03:00:00-0500", NY),
    ("2012-11-04T03:00:00-0500", "0 0 3 * * ?", "2012-11-05T03:00:00-0500", NY),
    # Unsatisfiable
    (J + "23:35 2012", "0 0 0 30 Feb ?", "", "UTC"),
    (J + "23:35 2012", "0 0 0 31 Apr ?", "", "UTC"),
    # Monthly job
    ("2012-11-04T00:00:00-0400", "0 0 3 3 * ?", "2012-12-03T03:00:00-0500", NY),
    # DST resulting in midnight not being a valid time (robfig/cron issue 157)
    ("2018-10-17T05:00:00-0400", "TZ=America/Sao_Paulo 0 0 9 10 * ?", "2018-11-10T06:00:00-0500", SP),
    ("2018-02-14T05:00:00-0500", "TZ=America/Sao_Paulo 0 0 9 22 * ?", "2018-02-22T07:00:00-0500", SP),
]

ACTIVATION = [  # (time, 5-field spec, expected): expected == (Next(time - 1s) == time)
    # Every fifteen minutes.
    (J + "15:00 2012", "0/15 * * * *", True),
    (J + "15:45 2012", "0/15 * * * *", True),
    (J + "15:40 2012", "0/15 * * * *", False),
    # Every fifteen minutes, starting at 5 minutes.
    (J + "15:05 2012", "5/15 * * * *", True),
    (J + "15:20 2012", "5/15 * * * *", True),
    (J + "15:50 2012", "5/15 * * * *", True),
    # Named months
    ("Sun Jul 15 15:00 2012", "0/15 * * Jul *", True),
    ("Sun Jul 15 15:00 2012", "0/15 * * Jun *", False),
    # Everything set.
    ("Sun Jul 15 08:30 2012", "30 08 ? Jul Sun", True),
    ("Sun Jul 15 08:30 2012", "30 08 15 Jul ?", True),
    ("Mon Jul 16 08:30 2012", "30 08 ? Jul Sun", False),
    ("Mon Jul 16 08:30 2012", "30 08 15 Jul ?", False),
    # Predefined schedules
    (J + "15:00 2012", "@hourly", True),
    (J + "15:04 2012", "@hourly", False),
    (J + "15:00 2012", "@daily", False),
    (J + "00:00 2012", "@daily", True),
    (J + "00:00 2012", "@weekly", False),
    ("Sun Jul 8 00:00 2012", "@weekly", True),
    ("Sun Jul 8 01:00 2012", "@weekly", False),
    ("Sun Jul 8 00:00 2012", "@monthly", False),
    ("Sun Jul 1 00:00 2012", "@monthly", True),
    # Interaction of DOW and DOM: if both are restricted, only one needs to match.
    ("Sun Jul 15 00:00 2012", "* * 1,15 * Sun", True),
    ("Fri Jun 15 00:00 2012", "* * 1,15 * Sun", True),
    ("Wed Aug 1 00:00 2012", "* * 1,15 * Sun", True),
    ("Sun Jul 15 00:00 2012", "* * */10 * Sun", True),  # upstream: "verifies #70"
    # However, if one has a star, then both need to match.
    ("Sun Jul 15 00:00 2012", "* * * * Mon", False),
    (J + "00:00 2012", "* * 1,15 * *", False),
    ("Sun Jul 15 00:00 2012", "* * 1,15 * *", True),
    ("Sun Jul 15 00:00 2012", "* * */2 * Sun", True),
]

STAR = 1 << 63
ALL = {"second": ((1 << 60) - 1) | STAR, "minute": ((1 << 60) - 1) | STAR, "hour": ((1 << 24) - 1) | STAR,
       "dom": ((1 << 32) - 2) | STAR, "month": ((1 << 13) - 2) | STAR, "dow": ((1 << 7) - 1) | STAR}

PARSE_OK = [  # (parser, spec, expected fields that the upstream row pins; the rest are all(field))
    ("standard", "5 * * * *", {"second": 1, "minute": 1 << 5}),                       # TestStandardSpecSchedule
    ("seconds", "0 5 * * * *", {"second": 1, "minute": 1 << 5}),                      # TestParseSchedule every5min
    ("seconds", "* 5 * * * *", {"minute": 1 << 5}),
    ("seconds", "@midnight", {"second": 1, "minute": 1, "hour": 1}),
    ("seconds", "@yearly", {"second": 1, "minute": 1, "hour": 1, "dom": 1 << 1, "month": 1 << 1}),
    ("seconds", "@annually", {"second": 1, "minute": 1, "hour": 1, "dom": 1 << 1, "month": 1 << 1}),
    # TestRange rows, placed in a field with the bounds upstream's row uses where one exists
    ("standard", "5-7/2 * * * *", {"second": 1, "minute": 1 << 5 | 1 << 7}),
    ("standard", "5-7/1 * * * *", {"second": 1, "minute": 1 << 5 | 1 << 6 | 1 << 7}),
    ("standard", "5-5 * * * *", {"second": 1, "minute": 1 << 5}),
    ("standard", "5-6/2 * * * *", {"second": 1, "minute": 1 << 5}),
]
PARSE_EVERY = [("standard", "@every 5m", 300), ("seconds", "@every 5m", 300)]
PARSE_ERR = [  # (parser, spec, substring of upstream's error)
    ("seconds", "* 5 j * * *", "failed to parse int from"),       # TestParseScheduleErrors
    ("seconds", "@every Xm", "failed to parse duration"),
    ("seconds", "@unrecognized", "unrecognized descriptor"),
    ("seconds", "", "empty spec string"),
    ("standard", "5 j * * *", "failed to parse int from"),        # TestStandardSpecSchedule
    ("standard", "* * * *", "expected exactly 5 fields"),
    ("standard", "5--5 * * * *", "too many hyphens"),             # TestRange error rows
    ("standard", "jan-x * * * *", "failed to parse int from"),
    ("standard", "2-x * * * *", "failed to parse int from"),
    ("standard", "*/-12 * * * *", "negative number"),
    ("standard", "*//2 * * * *", "too many slashes"),
    ("standard", "* * 0 * *", "below minimum"),
    ("standard", "60 * * * *", "above maximum"),
    ("standard", "5-3 * * * *", "beyond end of range"),
    ("standard", "*/0 * * * *", "should be a positive number"),
]


def main():
    out = {
        "provenance": "[UPSTREAM-RECALLED] robfig/cron v3.0.1 spec_test.go / parser_test.go rows written from knowledge "
                      "of the public upstream tests; not a Go run, not generated from this repo's oracle",
        "next": [dict(time=t, spec=s, expected=e, zone=z, **{"from": epoch(t)}, want=(epoch(e) if e else None))
                 for t, s, e, z in NEXT],
        "activation": [dict(time=t, spec=s, at=epoch(t), expected=e) for t, s, e in ACTIVATION],
        "parse_ok": [dict(parser=p, spec=s, masks=[str({**ALL, **f}[k]) for k in ("second", "minute", "hour", "dom", "month", "dow")])
                     for p, s, f in PARSE_OK],
        "parse_every": [dict(parser=p, spec=s, every_sec=v) for p, s, v in PARSE_EVERY],
        "parse_err": [dict(parser=p, spec=s, err=e) for p, s, e in PARSE_ERR],
    }
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "upstream_robfig.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1)
        f.write("\n")
    print(path, {k: len(v) for k, v in out.items() if isinstance(v, list)})


if __name__ == "__main__":
    main()
