// Package main — cgo driver of the B200 node validator.
//
// NOT COMPILED IN THIS ENVIRONMENT: the build image and the GPU box have no Go toolchain
// (`go: command not found`).  This file mirrors host/validator_core.cpp function for function over
// the same C ABI (include/b200validate.h); the C++ twin is what the tests build and run.  Keep the
// two in step: a behaviour change goes into both.
//
// Contract (operator-validator, SURVEY.md §8b): COMPONENT / WITH_WAIT / WITH_WORKLOAD /
// OUTPUT_DIR environment (each mirrored by a flag), exit code 0 or 1, empty status files
// <OUTPUT_DIR>/{cuda,driver,toolkit}-ready created on success and removed before validating.
//
// Build (on a machine that has Go):
//   CGO_CFLAGS="-I${REPO}/include" CGO_LDFLAGS="-L${REPO}/nvidia-gpu-operator-k8s-cluster_b200/lib -lb200validate" \
//   go build -o nvidia-validator ./go
package main

/*
#include <stdlib.h>
#include "b200validate.h"
*/
import "C"

import (
	"flag"
	"fmt"
	"io"
	"math"
	"net/http"
	"os"
	"os/exec"
	"path/filepath"
	"runtime"
	"strconv"
	"strings"
	"time"
	"unsafe"
)

type config struct {
	component, outputDir, nodeName, namespace, mode, nvidiaSmi, driverRoot string
	withWait, withWorkload                                                bool
	elems                                                                 uint64
	gpus, sleepSeconds, maxRetries, childDevice                           int
	injectFault                                                           int64
}

func envOr(name, dflt string) string {
	if v := os.Getenv(name); v != "" {
		return v
	}
	return dflt
}

func envInt(name string, dflt int) int {
	if v, err := strconv.Atoi(os.Getenv(name)); err == nil {
		return v
	}
	return dflt
}

func envBool(name string, dflt bool) bool {
	if v, err := strconv.ParseBool(os.Getenv(name)); err == nil {
		return v
	}
	return dflt
}

// parseConfig: environment first, flags win (same order as parse_config in the C++ twin).
func parseConfig() (*config, error) {
	c := &config{}
	fs := flag.NewFlagSet("nvidia-validator", flag.ContinueOnError)
	fs.StringVar(&c.component, "component", envOr("COMPONENT", ""), "component to validate: cuda, driver, toolkit")
	fs.BoolVar(&c.withWait, "with-wait", envBool("WITH_WAIT", false), "retry until the validation succeeds")
	fs.BoolVar(&c.withWorkload, "with-workload", envBool("WITH_WORKLOAD", true), "run the CUDA workload")
	fs.StringVar(&c.outputDir, "output-dir", envOr("OUTPUT_DIR", "/run/nvidia/validations"), "status file directory")
	fs.StringVar(&c.nodeName, "node-name", envOr("NODE_NAME", ""), "node name")
	fs.StringVar(&c.namespace, "namespace", envOr("OPERATOR_NAMESPACE", ""), "operator namespace")
	fs.StringVar(&c.mode, "mode", envOr("B200V_MODE", "validate"), "validate | stream | nccl")
	fs.Uint64Var(&c.elems, "elems", uint64(envInt("B200V_ELEMS", 50000)), "elements per vector")
	fs.IntVar(&c.gpus, "gpus", envInt("B200V_GPUS", 0), "devices to validate (0 = all visible)")
	if err := fs.Parse(os.Args[1:]); err != nil {
		return nil, err
	}
	c.sleepSeconds = envInt("B200V_SLEEP_SECONDS", 5)
	c.maxRetries = envInt("B200V_MAX_RETRIES", 0)
	c.childDevice = envInt("B200V_CHILD_DEVICE", -1)
	c.injectFault = int64(envInt("B200V_INJECT_FAULT", -1))
	c.nvidiaSmi = envOr("B200V_NVIDIA_SMI", "nvidia-smi")
	c.driverRoot = envOr("DRIVER_ROOT", "/run/nvidia/driver")
	if c.mode != "validate" && c.mode != "stream" && c.mode != "nccl" && c.mode != "nvls" {
		return nil, fmt.Errorf("B200V_MODE must be validate, stream, nccl or nvls, not %s", c.mode)
	}
	return c, nil
}

func strerror(code C.int) string { return C.GoString(C.b200v_strerror(code)) }

func statusPath(c *config, name string) string { return filepath.Join(c.outputDir, name) }

func createStatusFile(c *config, name string) error {
	f, err := os.Create(statusPath(c, name))
	if err != nil {
		return err
	}
	return f.Close()
}

func deleteStatusFile(c *config, name string) error {
	err := os.Remove(statusPath(c, name))
	if err != nil && !os.IsNotExist(err) {
		return err
	}
	return nil
}

// glibcRand reproduces the sample's unseeded rand() stream (TYPE_3, seed 1) so the Go driver fills
// the same inputs the C sample does without calling into libc from Go.
type glibcRand struct {
	r []uint32
	k int
}

func newGlibcRand(seed int32) *glibcRand {
	g := &glibcRand{r: make([]uint32, 34, 4096)}
	g.r[0] = uint32(seed)
	for i := 1; i < 31; i++ {
		hi, lo := int64(int32(g.r[i-1]))/127773, int64(int32(g.r[i-1]))%127773
		w := 16807*lo - 2836*hi
		if w < 0 {
			w += 2147483647
		}
		g.r[i] = uint32(w)
	}
	for i := 31; i < 34; i++ {
		g.r[i] = g.r[i-31]
	}
	for i := 34; i < 344; i++ {
		g.r = append(g.r, g.r[i-31]+g.r[i-3])
	}
	g.k = 344
	return g
}

func (g *glibcRand) next() int32 {
	v := g.r[g.k-31] + g.r[g.k-3]
	g.r = append(g.r, v)
	g.k++
	return int32(v >> 1)
}

// runVectorAddSample mirrors run_vectoradd_sample: host init, device add through the C ABI, the
// sample's verification loop, then a bitwise check.
func runVectorAddSample(dev int, n uint64, injectFault int64, timingOut *C.b200v_timing, beforeDevice func()) int {
	runtime.LockOSThread() // CUDA's current device is per OS thread
	defer runtime.UnlockOSThread()
	fmt.Printf("[Vector addition of %d elements]\n", n)
	a, b, out := make([]float32, n+1), make([]float32, n+1), make([]float32, n+1)
	rng := newGlibcRand(1)
	const randMax = float32(2147483647)
	for i := uint64(0); i < n; i++ {
		a[i] = float32(rng.next()) / randMax
		b[i] = float32(rng.next()) / randMax
	}
	if beforeDevice != nil {
		beforeDevice() // fan-out children meet here: host init done, GPU work next
	}
	fmt.Println("Copy input data from the host memory to the CUDA device")
	var t C.b200v_timing
	rc := C.b200v_vectoradd(C.int(dev), (*C.float)(unsafe.Pointer(&a[0])), (*C.float)(unsafe.Pointer(&b[0])),
		(*C.float)(unsafe.Pointer(&out[0])), C.size_t(n), &t)
	if rc != 0 {
		fmt.Fprintf(os.Stderr, "Failed to launch vectorAdd kernel (error code %s)!\n", strerror(rc))
		return 1
	}
	if timingOut != nil {
		*timingOut = t
	}
	fmt.Printf("CUDA kernel launch with %d blocks of %d threads\n", int(t.grid), int(t.block))
	fmt.Println("Copy output data from the CUDA device to the host memory")
	if injectFault >= 0 && uint64(injectFault) < n {
		out[injectFault] += 1.0
	}
	for i := uint64(0); i < n; i++ {
		if math.Abs(float64(a[i]+b[i]-out[i])) > 1e-5 {
			fmt.Fprintf(os.Stderr, "Result verification failed at element %d!\n", i)
			return 1
		}
	}
	for i := uint64(0); i < n; i++ {
		want := a[i] + b[i] + 0.0
		if math.Float32bits(want) != math.Float32bits(out[i]) {
			fmt.Fprintf(os.Stderr, "Result verification failed at element %d! (bitwise)\n", i)
			return 1
		}
	}
	fmt.Println("Test PASSED")
	fmt.Println("Done")
	return 0
}

// fanOut mirrors fan_out: one child process per device (exit-code isolation), held at a start
// barrier until all are ready, each handing back one JSON row.  The child finds its pipe ends in
// B200V_FANOUT_FDS="ready_w,go_r,result_w" (ExtraFiles puts them at descriptors 3, 4, 5).
func fanOut(c *config, use int, args []string) ([]int, []string) {
	results, rows := make([]int, use), make([]string, use)
	self, _ := os.Executable()
	readyR, readyW, _ := os.Pipe()
	goR, goW, _ := os.Pipe()
	cmds := make([]*exec.Cmd, use)
	resR := make([]*os.File, use)
	for i := 0; i < use; i++ {
		r, w, _ := os.Pipe()
		cmd := exec.Command(self, args...)
		cmd.Env = append(os.Environ(), "B200V_CHILD_DEVICE="+strconv.Itoa(i), "B200V_FANOUT_FDS=3,4,5")
		cmd.ExtraFiles = []*os.File{readyW, goR, w}
		cmd.Stdout, cmd.Stderr = os.Stdout, os.Stderr
		if err := cmd.Start(); err != nil {
			results[i] = 1
		} else {
			cmds[i] = cmd
		}
		w.Close()
		resR[i] = r
	}
	readyW.Close()
	goR.Close()
	limit := time.Duration(envInt("B200V_FANOUT_BARRIER_TIMEOUT_S", 300)) * time.Second
	readyR.SetReadDeadline(time.Now().Add(limit))
	buf := make([]byte, 64)
	for arrived := 0; arrived < use; {
		n, err := readyR.Read(buf)
		if n <= 0 || err != nil {
			break // every writer gone, or the timeout: release whoever is waiting
		}
		arrived += n
	}
	goW.Close()
	readyR.Close()
	for i, cmd := range cmds {
		if cmd != nil {
			if err := cmd.Wait(); err != nil {
				results[i] = 1
			}
		}
		b, _ := io.ReadAll(resR[i])
		resR[i].Close()
		rows[i] = string(b)
	}
	return results, rows
}

// fanoutChildMain mirrors fanout_child_main: one device, no status files, one result row.
func fanoutChildMain(c *config) int {
	var readyW, goR, resW *os.File
	var a, b, d int
	if n, _ := fmt.Sscanf(envOr("B200V_FANOUT_FDS", ""), "%d,%d,%d", &a, &b, &d); n == 3 {
		readyW, goR, resW = os.NewFile(uintptr(a), "ready"), os.NewFile(uintptr(b), "go"), os.NewFile(uintptr(d), "result")
	}
	dev := c.childDevice
	inj := int64(-1)
	if d := envInt("B200V_INJECT_DEVICE", -1); d < 0 || d == dev {
		inj = c.injectFault
	}
	var warm [48]float32 // context and streams exist before the barrier
	C.b200v_vectoradd(C.int(dev), (*C.float)(unsafe.Pointer(&warm[0])), (*C.float)(unsafe.Pointer(&warm[16])),
		(*C.float)(unsafe.Pointer(&warm[32])), 16, nil)
	meet := func() {
		if readyW != nil {
			readyW.Write([]byte{1})
			readyW.Close()
		}
		if goR != nil {
			io.ReadAll(goR) // returns when the parent closes its end
			goR.Close()
		}
	}
	var t C.b200v_timing
	rc := runVectorAddSample(dev, c.elems, inj, &t, meet)
	bytes := 12.0 * float64(c.elems)
	gbs := func(ms float64) float64 {
		if ms <= 0 {
			return 0
		}
		return bytes / (ms * 1e-3) / 1e9
	}
	passed := 0
	if rc == 0 {
		passed = 1
	}
	row := fmt.Sprintf("{\"mode\":\"fanout-child\",\"gpu\":%d,\"elems\":%d,\"passed\":%d,\"h2d_ms\":%.4f,"+
		"\"kernel_ms\":%.4f,\"d2h_ms\":%.4f,\"total_ms\":%.4f,\"kernel_launches\":%d,\"kernel_gbs\":%.1f,\"e2e_gbs\":%.2f",
		dev, c.elems, passed, float64(t.h2d_ms), float64(t.kernel_ms), float64(t.d2h_ms), float64(t.total_ms),
		int(t.kernel_launches), gbs(float64(t.kernel_ms)), gbs(float64(t.total_ms)))
	hbmLog2 := envInt("B200V_FANOUT_HBM_LOG2", -1)
	if hbmLog2 < 0 {
		hbmLog2 = 0
		if c.elems >= 1<<22 {
			for uint64(2)<<uint(hbmLog2) <= c.elems {
				hbmLog2++
			}
		}
	}
	rc2 := 0
	if rc == 0 && hbmLog2 > 0 && hbmLog2 <= 32 {
		var r C.b200v_stream_result
		src := C.b200v_stream(C.int(dev), C.B200V_OP_VECTORADD, C.size_t(1)<<uint(hbmLog2), 0,
			C.uint64_t(0xB200+0x1000*uint64(dev)), 5, 20, &r)
		if src != 0 || r.mismatches != 0 {
			fmt.Fprintf(os.Stderr, "device-resident vectorAdd on gpu %d failed: %s\n", dev, strerror(src))
			rc2 = 1
		}
		row += fmt.Sprintf(",\"hbm_log2_n\":%d,\"hbm_median_ms\":%.5f,\"hbm_median_gbs\":%.1f,\"hbm_best_gbs\":%.1f,"+
			"\"hbm_mismatches\":%d,\"hbm_digest\":\"%016x\"", hbmLog2, float64(r.median_ms), float64(r.median_gbs),
			float64(r.best_gbs), uint64(r.mismatches), uint64(r.digest))
	}
	if resW != nil {
		resW.WriteString(row + "}\n")
		resW.Close()
	}
	if rc == 0 && rc2 == 0 {
		return 0
	}
	return 1
}

// reportFanout mirrors report_fanout: the children's rows, then the whole-node line and the
// fanout-bandwidth.prom file (B200V_WRITE_PROM).
func reportFanout(c *config, use int, rows []string) {
	num := func(row, key string, dflt float64) float64 {
		k := "\"" + key + "\":"
		at := strings.Index(row, k)
		if at < 0 {
			return dflt
		}
		rest := row[at+len(k):]
		end := strings.IndexAny(rest, ",}")
		if end < 0 {
			end = len(rest)
		}
		if v, err := strconv.ParseFloat(strings.Trim(rest[:end], "\""), 64); err == nil {
			return v
		}
		return dflt
	}
	var perHbm, perE2e []string
	var sumHbm, slowHbm, slowTotal, hbmBytes float64
	prom := "# HELP b200v_fanout_hbm_bandwidth_gbs device-resident vectorAdd GB/s of each GPU, all GPUs running at once\n" +
		"# TYPE b200v_fanout_hbm_bandwidth_gbs gauge\n"
	reported := 0
	for i, row := range rows {
		if row == "" {
			continue
		}
		reported++
		fmt.Print(row)
		if hbm := num(row, "hbm_median_gbs", -1); hbm >= 0 {
			sumHbm += hbm
			slowHbm = math.Max(slowHbm, num(row, "hbm_median_ms", 0))
			hbmBytes = 12.0 * math.Ldexp(1, int(num(row, "hbm_log2_n", 0)))
			perHbm = append(perHbm, fmt.Sprintf("%.1f", hbm))
			prom += fmt.Sprintf("b200v_fanout_hbm_bandwidth_gbs{gpu=\"%d\"} %.1f\n", i, hbm)
		}
		slowTotal = math.Max(slowTotal, num(row, "total_ms", 0))
		perE2e = append(perE2e, fmt.Sprintf("%.2f", num(row, "e2e_gbs", 0)))
	}
	if reported == 0 {
		return
	}
	jobHbm, jobE2e := 0.0, 0.0
	if slowHbm > 0 {
		jobHbm = float64(len(perHbm)) * hbmBytes / (slowHbm * 1e-3) / 1e9
	}
	if slowTotal > 0 {
		jobE2e = float64(reported) * 12.0 * float64(c.elems) / (slowTotal * 1e-3) / 1e9
	}
	fmt.Printf("{\"mode\":\"fanout\",\"gpus\":%d,\"reported\":%d,\"elems\":%d,\"per_gpu_hbm_gbs\":[%s],"+
		"\"sum_hbm_gbs\":%.1f,\"aggregate_hbm_gbs\":%.1f,\"slowest_hbm_ms\":%.5f,\"per_gpu_e2e_gbs\":[%s],"+
		"\"aggregate_e2e_gbs\":%.2f,\"slowest_total_ms\":%.4f}\n", use, reported, c.elems, strings.Join(perHbm, ","),
		sumHbm, jobHbm, slowHbm, strings.Join(perE2e, ","), jobE2e, slowTotal)
	if len(perHbm) > 0 && envBool("B200V_WRITE_PROM", false) {
		prom += fmt.Sprintf("# HELP b200v_fanout_hbm_aggregate_gbs all GPUs' bytes over the slowest GPU's time\n"+
			"# TYPE b200v_fanout_hbm_aggregate_gbs gauge\nb200v_fanout_hbm_aggregate_gbs %.1f\n", jobHbm)
		os.WriteFile(statusPath(c, "fanout-bandwidth.prom"), []byte(prom), 0o644)
	}
}

// validateCuda mirrors validate_cuda: one validator instance per device, status files per device,
// cuda-ready only when every device passed.
func validateCuda(c *config) int {
	if err := deleteStatusFile(c, "cuda-ready"); err != nil {
		fmt.Fprintf(os.Stderr, "unable to remove stale status files under %s: %v\n", c.outputDir, err)
		return 1
	}
	old, _ := filepath.Glob(statusPath(c, "cuda-gpu*-ready"))
	for _, f := range old {
		os.Remove(f)
	}
	if !c.withWorkload { // upstream: no workload, just (re)create the status file
		if err := createStatusFile(c, "cuda-ready"); err != nil {
			fmt.Fprintf(os.Stderr, "unable to create status file %s: %v\n", statusPath(c, "cuda-ready"), err)
			return 1
		}
		fmt.Println("cuda workload validation is disabled (WITH_WORKLOAD=false)")
		return 0
	}
	var ndev C.int
	if rc := C.b200v_device_count(&ndev); rc != 0 || ndev < 1 {
		fmt.Fprintf(os.Stderr, "cuda validation failed: %s\n", strerror(rc))
		return 1
	}
	if c.gpus > int(ndev) {
		fmt.Fprintf(os.Stderr, "cuda validation failed: B200V_GPUS=%d but only %d device(s) visible\n", c.gpus, int(ndev))
		return 1
	}
	use := int(ndev)
	if c.gpus > 0 {
		use = c.gpus
	}
	results := make([]int, use)
	if use == 1 && !envBool("B200V_FANOUT_FORCE", false) { // one device: in this process, as upstream
		results[0] = runVectorAddSample(0, c.elems, c.injectFault, nil, nil)
	} else {
		var rows []string
		results, rows = fanOut(c, use, []string{"--component", "cuda", "--output-dir", c.outputDir, "--elems",
			strconv.FormatUint(c.elems, 10)})
		reportFanout(c, use, rows)
	}
	all := true
	for i, rc := range results {
		if rc == 0 {
			if err := createStatusFile(c, fmt.Sprintf("cuda-gpu%d-ready", i)); err != nil {
				fmt.Fprintf(os.Stderr, "unable to create status file for gpu %d: %v\n", i, err)
				all = false
			}
		} else {
			fmt.Fprintf(os.Stderr, "cuda validation failed on gpu %d (exit code %d)\n", i, rc)
			all = false
		}
	}
	if !all {
		return 1
	}
	if err := createStatusFile(c, "cuda-ready"); err != nil {
		fmt.Fprintf(os.Stderr, "unable to create status file %s: %v\n", statusPath(c, "cuda-ready"), err)
		return 1
	}
	fmt.Printf("cuda workload validation is successful on %d device(s)\n", use)
	return 0
}

func validateWithSmi(c *config, file, what string, argv []string) int {
	if err := deleteStatusFile(c, file); err != nil {
		fmt.Fprintf(os.Stderr, "unable to remove stale %s: %v\n", file, err)
		return 1
	}
	cmd := exec.Command(argv[0], argv[1:]...)
	cmd.Stdout, cmd.Stderr = os.Stdout, os.Stderr
	if err := cmd.Run(); err != nil {
		fmt.Fprintf(os.Stderr, "%s validation failed: `%s`: %v\n", what, strings.Join(argv, " "), err)
		return 1
	}
	if err := createStatusFile(c, file); err != nil {
		fmt.Fprintf(os.Stderr, "unable to create status file %s: %v\n", file, err)
		return 1
	}
	fmt.Printf("%s validation is successful\n", what)
	return 0
}

func validateDriver(c *config) int {
	argv := []string{c.nvidiaSmi}
	if _, err := os.Stat(filepath.Join(c.driverRoot, "usr/bin/nvidia-smi")); err == nil && c.nvidiaSmi == "nvidia-smi" {
		argv = []string{"chroot", c.driverRoot, "nvidia-smi"}
	}
	return validateWithSmi(c, "driver-ready", "driver", argv)
}

func validateToolkit(c *config) int {
	return validateWithSmi(c, "toolkit-ready", "toolkit", []string{c.nvidiaSmi})
}

// readAllocatableGPUs mirrors read_allocatable_gpus: B200V_ALLOCATABLE_GPUS, else the first
// integer after "nvidia.com/gpu" (or a bare number) in B200V_ALLOCATABLE_FILE; -1 = no source.
func readAllocatableGPUs() int64 {
	if v := os.Getenv("B200V_ALLOCATABLE_GPUS"); v != "" {
		n, err := strconv.ParseInt(v, 10, 64)
		if err != nil || n < 0 {
			return 0
		}
		return n
	}
	path := os.Getenv("B200V_ALLOCATABLE_FILE")
	if path == "" {
		return -1
	}
	raw, err := os.ReadFile(path)
	if err != nil {
		return 0
	}
	text := strings.TrimSpace(string(raw))
	if n, err := strconv.ParseInt(text, 10, 64); err == nil && n >= 0 {
		return n
	}
	at := strings.Index(text, "nvidia.com/gpu")
	if at < 0 {
		return 0
	}
	rest := strings.TrimLeft(text[at+len("nvidia.com/gpu"):], ":\"'= \t")
	end := 0
	for end < len(rest) && rest[end] >= '0' && rest[end] <= '9' {
		end++
	}
	n, err := strconv.ParseInt(rest[:end], 10, 64)
	if err != nil {
		return 0
	}
	return n
}

// validatePlugin mirrors validate_plugin (SURVEY.md §8f-1).
func validatePlugin(c *config) int {
	if err := deleteStatusFile(c, "plugin-ready"); err != nil {
		fmt.Fprintf(os.Stderr, "unable to remove stale plugin-ready: %v\n", err)
		return 1
	}
	gpus := readAllocatableGPUs()
	if gpus == -1 {
		fmt.Fprintln(os.Stderr, "plugin validation needs the node's allocatable nvidia.com/gpu count: set "+
			"B200V_ALLOCATABLE_GPUS or B200V_ALLOCATABLE_FILE (no Kubernetes client in this build)")
		return 1
	}
	retries := envInt("B200V_PLUGIN_RETRIES", 30)
	for attempt := 1; gpus <= 0 && attempt < retries; attempt++ {
		fmt.Fprintf(os.Stderr, "GPU resources are not yet discovered by the node, retry: %d\n", attempt)
		time.Sleep(time.Duration(c.sleepSeconds) * time.Second)
		gpus = readAllocatableGPUs()
	}
	if gpus <= 0 {
		fmt.Fprintln(os.Stderr, "GPU resources are not discovered by the node")
		return 1
	}
	fmt.Printf("node advertises nvidia.com/gpu: %d\n", gpus)
	if c.withWorkload && runVectorAddSample(0, c.elems, c.injectFault, nil, nil) != 0 {
		return 1
	}
	if err := createStatusFile(c, "plugin-ready"); err != nil {
		fmt.Fprintf(os.Stderr, "unable to create status file plugin-ready: %v\n", err)
		return 1
	}
	// upstream's file is empty; the advertised count is kept in it for the metrics mode
	os.WriteFile(statusPath(c, "plugin-ready"), []byte(fmt.Sprintf("nvidia.com/gpu: %d\n", gpus)), 0o644)
	fmt.Println("plugin validation is successful")
	return 0
}

// renderMetrics mirrors render_metrics (SURVEY.md §8f-2): status-file gauges plus *.prom files.
func renderMetrics(c *config) string {
	var b strings.Builder
	for _, comp := range []string{"driver", "toolkit", "cuda", "plugin"} {
		name := "gpu_operator_node_" + comp + "_ready"
		v, ts := 0, int64(0)
		if fi, err := os.Stat(statusPath(c, comp+"-ready")); err == nil {
			v, ts = 1, fi.ModTime().Unix()
		}
		fmt.Fprintf(&b, "# HELP %s 1 when %s validation has left its status file on this node\n# TYPE %s gauge\n%s %d\n",
			name, comp, name, name, v)
		tsName := "gpu_operator_node_" + comp + "_validation_last_success_ts_seconds"
		fmt.Fprintf(&b, "# HELP %s unix time of the last successful %s validation (0 = never)\n# TYPE %s gauge\n%s %d\n",
			tsName, comp, tsName, tsName, ts)
	}
	perGPU, _ := filepath.Glob(statusPath(c, "cuda-gpu*-ready"))
	b.WriteString("# HELP b200v_node_cuda_gpu_ready 1 for every GPU whose own vectorAdd validation passed\n# TYPE b200v_node_cuda_gpu_ready gauge\n")
	for _, f := range perGPU {
		idx := strings.TrimSuffix(strings.TrimPrefix(filepath.Base(f), "cuda-gpu"), "-ready")
		fmt.Fprintf(&b, "b200v_node_cuda_gpu_ready{gpu=\"%s\"} 1\n", idx)
	}
	advertised := int64(0) // what plugin validation read from the node's allocatable nvidia.com/gpu
	if raw, err := os.ReadFile(statusPath(c, "plugin-ready")); err == nil {
		if at := strings.Index(string(raw), "nvidia.com/gpu:"); at >= 0 {
			if v, err := strconv.ParseInt(strings.TrimSpace(string(raw)[at+len("nvidia.com/gpu:"):]), 10, 64); err == nil && v > 0 {
				advertised = v
			}
		}
	}
	fmt.Fprintf(&b, "# HELP gpu_operator_node_device_plugin_devices_total GPUs the device plugin advertises on this node (allocatable nvidia.com/gpu seen by plugin validation)\n"+
		"# TYPE gpu_operator_node_device_plugin_devices_total gauge\ngpu_operator_node_device_plugin_devices_total %d\n", advertised)
	proms, _ := filepath.Glob(statusPath(c, "*.prom"))
	for _, f := range proms {
		if raw, err := os.ReadFile(f); err == nil {
			b.Write(raw)
		}
	}
	return b.String()
}

func runStreamSweep(c *config) int {
	ops := []struct {
		id   C.int
		name string
	}{{C.B200V_OP_COPY, "copy"}, {C.B200V_OP_SCALE, "scale"}, {C.B200V_OP_ADD, "add"}, {C.B200V_OP_TRIAD, "triad"}}
	failures := 0
	if c.childDevice < 0 && c.gpus > 1 { // one sweep per device, each in its own process, all at once
		results, _ := fanOut(c, c.gpus, []string{"--mode", "stream"})
		for i, rc := range results {
			if rc != 0 {
				fmt.Fprintf(os.Stderr, "stream sweep failed on gpu %d (exit code %d)\n", i, rc)
				failures++
			}
		}
		if failures > 0 {
			return 1
		}
		return 0
	}
	dev := 0
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	if c.childDevice >= 0 { // per-GPU sweeps start together: create the context, then meet the others
		dev = c.childDevice
		var warm C.b200v_info
		C.b200v_device_info(C.int(dev), &warm)
		var a, b, d int
		if n, _ := fmt.Sscanf(envOr("B200V_FANOUT_FDS", ""), "%d,%d,%d", &a, &b, &d); n == 3 {
			readyW, goR := os.NewFile(uintptr(a), "ready"), os.NewFile(uintptr(b), "go")
			readyW.Write([]byte{1})
			readyW.Close()
			io.ReadAll(goR)
			goR.Close()
		}
	}
	for lg := envInt("B200V_MIN_LOG2", 18); lg <= envInt("B200V_MAX_LOG2", 32); lg++ {
		for _, op := range ops {
			var r C.b200v_stream_result
			rc := C.b200v_stream(C.int(dev), op.id, C.size_t(1)<<uint(lg), 3.0, 0xB200, 5, 20, &r)
			if rc != 0 && rc != C.B200V_E_MISMATCH {
				fmt.Fprintf(os.Stderr, "stream %s n=2^%d failed: %s\n", op.name, lg, strerror(rc))
				failures++
				continue
			}
			if r.mismatches != 0 {
				failures++
			}
			fmt.Printf("{\"mode\":\"stream\",\"gpu\":%d,\"op\":\"%s\",\"log2_n\":%d,\"median_gbs\":%.1f,\"best_gbs\":%.1f,"+
				"\"l2_resident\":%d,\"mismatches\":%d,\"digest\":\"%016x\"}\n", dev, op.name, lg, float64(r.median_gbs),
				float64(r.best_gbs), int(r.l2_resident), uint64(r.mismatches), uint64(r.digest))
		}
	}
	if failures > 0 {
		return 1
	}
	return 0
}

func runNcclProbe(c *config) int {
	var ndev C.int
	if rc := C.b200v_device_count(&ndev); rc != 0 || ndev < 1 {
		fmt.Fprintf(os.Stderr, "nccl probe failed: %s\n", strerror(rc))
		return 1
	}
	use := int(ndev)
	if c.gpus > 0 {
		use = c.gpus
	}
	devs := make([]C.int, use)
	for i := range devs {
		devs[i] = C.int(i)
	}
	rows := make([]C.b200v_ar_row, 64)
	var nrows C.int
	lo, hi := C.size_t(1)<<uint(envInt("B200V_MIN_LOG2", 10)), C.size_t(1)<<uint(envInt("B200V_MAX_LOG2", 30))
	var rc C.int
	if c.mode == "nvls" { // the hand-written NVLink kernels: same sweep, sizes multiple of 16 bytes
		if lo < 16 {
			lo = 16
		}
		rc = C.b200v_nvls_allreduce_probe(C.int(use), &devs[0], lo, hi, 5, 20, &rows[0], 64, &nrows)
	} else {
		rc = C.b200v_allreduce_probe(C.int(use), &devs[0], lo, hi, 5, 20, &rows[0], 64, &nrows)
	}
	failures := 0
	if rc != 0 {
		fmt.Fprintf(os.Stderr, "%s probe failed: %s\n", c.mode, strerror(rc))
		failures++
	}
	for i := 0; i < int(nrows); i++ {
		r := rows[i]
		if r.mismatches != 0 {
			failures++
		}
		fmt.Printf("{\"mode\":\"%s\",\"ranks\":%d,\"bytes\":%d,\"busbw_gbs\":%.2f,\"frac_of_900\":%.4f,\"mismatches\":%d}\n",
			c.mode, use, uint64(r.bytes), float64(r.busbw_gbs), float64(r.busbw_gbs)/900.0, uint64(r.mismatches))
	}
	if failures > 0 {
		return 1
	}
	return 0
}

func main() {
	c, err := parseConfig()
	if err != nil {
		fmt.Fprintf(os.Stderr, "Incorrect Usage. %v\n", err)
		os.Exit(1)
	}
	if c.mode != "validate" { // a probe that never completes must not hang the node's validation
		if limit := envInt("B200V_PROBE_TIMEOUT_S", 600); limit > 0 {
			time.AfterFunc(time.Duration(limit)*time.Second, func() {
				fmt.Fprintln(os.Stderr, "probe timed out (B200V_PROBE_TIMEOUT_S)")
				os.Exit(1)
			})
		}
	}
	switch c.mode {
	case "stream":
		os.Exit(runStreamSweep(c))
	case "nccl", "nvls":
		os.Exit(runNcclProbe(c))
	}
	if c.component == "" {
		fmt.Fprintln(os.Stderr, "invalid -c <component-name> flag: must not be empty string")
		os.Exit(1)
	}
	if c.component == "metrics" { // node-status-exporter mode: serve /metrics on METRICS_PORT
		http.HandleFunc("/metrics", func(w http.ResponseWriter, _ *http.Request) {
			w.Header().Set("Content-Type", "text/plain; version=0.0.4")
			fmt.Fprint(w, renderMetrics(c))
		})
		fmt.Fprintln(os.Stderr, http.ListenAndServe(":"+envOr("METRICS_PORT", "8000"), nil))
		os.Exit(1)
	}
	if c.childDevice >= 0 {
		os.Exit(fanoutChildMain(c))
	}
	for tries := 0; ; {
		rc := -1
		switch c.component {
		case "cuda":
			rc = validateCuda(c)
		case "driver":
			rc = validateDriver(c)
		case "toolkit":
			rc = validateToolkit(c)
		case "plugin":
			rc = validatePlugin(c)
		}
		if rc == -1 {
			fmt.Fprintf(os.Stderr, "invalid component specified for validation: %s (this build supports cuda, driver, toolkit)\n", c.component)
			os.Exit(1)
		}
		if rc == 0 {
			os.Exit(0)
		}
		tries++
		if !c.withWait || (c.maxRetries > 0 && tries >= c.maxRetries) {
			os.Exit(1)
		}
		fmt.Fprintf(os.Stderr, "%s is not ready, retrying after %d seconds\n", c.component, c.sleepSeconds)
		time.Sleep(time.Duration(c.sleepSeconds) * time.Second)
	}
}
