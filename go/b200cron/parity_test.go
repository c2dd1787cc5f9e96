// parity_test.go closes the "parity unpinned" gap for anyone with a Go toolchain and network access: it runs the
// REAL robfig/cron v3.0.1 on fixed specs/zones/times and prints the answers in the layout of
// tests/golden/vectors.json ("next" section), so they can be diffed against the committed oracle vectors.
//
//	go mod init parity && go get github.com/robfig/cron/v3@v3.0.1 && go test -run TestDump -v
//
// UNVERIFIED: never compiled here (no Go toolchain on the build image).
package b200cron

import (
	"encoding/json"
	"os"
	"strings"
	"testing"
	"time"

	cronv3 "github.com/robfig/cron/v3"
)

type nextCase struct {
	Spec        string    `json:"spec"`
	WithSeconds bool      `json:"with_seconds"`
	Zone        string    `json:"zone"`
	From        []int64   `json:"from"`
	Want        [][]int64 `json:"want"`
}

func TestDump(t *testing.T) {
	raw, err := os.ReadFile("../../tests/golden/vectors.json")
	if err != nil {
		t.Skip(err)
	}
	var v struct {
		Next []nextCase `json:"next"`
	}
	if err := json.Unmarshal(raw, &v); err != nil {
		t.Fatal(err)
	}
	sec := cronv3.NewParser(cronv3.Second | cronv3.Minute | cronv3.Hour | cronv3.Dom | cronv3.Month | cronv3.Dow | cronv3.Descriptor)
	for _, c := range v.Next {
		loc, err := time.LoadLocation(c.Zone)
		if err != nil {
			t.Fatal(err)
		}
		var sched cronv3.Schedule
		if c.WithSeconds {
			sched, err = sec.Parse(c.Spec)
		} else {
			sched, err = cronv3.ParseStandard(c.Spec)
		}
		if err != nil {
			t.Fatal(err)
		}
		for i, f := range c.From {
			cur := time.Unix(f, 0).In(loc) // schedules without TZ= are evaluated in t's location
			for j := 0; j < 4; j++ {
				got := int64(-1 << 63)
				if !cur.IsZero() {
					cur = sched.Next(cur)
					if !cur.IsZero() {
						got = cur.Unix()
					}
				}
				if got != c.Want[i][j] {
					t.Errorf("%q %s from %d [%d]: robfig %d, oracle %d", c.Spec, c.Zone, f, j, got, c.Want[i][j])
				}
			}
		}
	}
}

// TestLocalZoneName: time.Local.String() is "Local" under a mounted /etc/localtime, so the shim must find the name
// itself; with TZ unset the answer comes from the /etc/localtime symlink (run on a host whose localtime is not UTC).
func TestLocalZoneName(t *testing.T) {
	t.Setenv("TZ", ":Europe/Berlin")
	if got := localZoneName(); got != "Europe/Berlin" {
		t.Fatalf("TZ=:Europe/Berlin -> %q", got)
	}
	t.Setenv("TZ", "")
	if got := localZoneName(); got != "UTC" {
		t.Fatalf("TZ= -> %q", got)
	}
	os.Unsetenv("TZ")
	if target, err := os.Readlink("/etc/localtime"); err == nil && strings.Contains(target, "zoneinfo/") {
		if got := localZoneName(); !strings.HasSuffix(target, got) || got == "" {
			t.Fatalf("/etc/localtime -> %q, got %q", target, got)
		}
	}
}
