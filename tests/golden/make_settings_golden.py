#!/usr/bin/env python
"""Golden fixtures for the command-line surface: runs the reference's own Settings class (oracle/_ref/ref_settings,
built by oracle/build_ref.sh from Source/Settings/Settings.cpp) on a list of command lines and stores
(argv, exit code, printed settings) in tests/golden/settings_cases.json."""
import json
import os
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
EXE = os.path.join(ROOT, "oracle", "_ref", "ref_settings")

CASES = [
    [],
    ["--3d", "--time-steps", "100", "--sizex", "80", "--same-size", "--use-tfsf", "--3d", "--angle-phi", "0", "--use-pml",
     "--dx", "0.0005", "--wavelength", "0.02", "--save-res"],                      # README launch example
    ["--2d", "--sizex", "4096", "--sizey", "4096", "--use-pml", "--time-steps", "1000"],
    ["--same-size", "--sizex", "64"],                                              # order sensitivity: copies 100, then x = 64
    ["--sizex", "64", "--same-size", "--pml-sizex", "7", "--same-size-pml", "--tfsf-sizex", "13", "--same-size-tfsf"],
    ["--tfsf-sizex", "9", "--same-size-ntff"],                                     # the typo: sets ntffSizeZ and tfsfSizeY
    ["--dx", "0.001", "--wavelength", "0.03", "--angle-teta", "60.5", "--angle-phi", "30", "--angle-psi", "45"],
    ["--log-level", "2", "--buffer-size", "3", "--num-cuda-gpus", "4", "--available-topologies", "topo.txt",
     "--topology-sizex", "2", "--topology-sizey", "3", "--topology-sizez", "4"],
    ["--use-ntff", "--use-metamaterials", "--use-amp-mode", "--amplitude-time-steps", "55", "--use-double-material-precision",
     "--parallel-grid", "--optimal-topology", "--save-materials", "--save-interm-res", "--interm-save-step", "7",
     "--save-scattered-field-res", "--save-scattered--field-interm", "--ntff-sizex", "3", "--ntff-sizey", "4", "--ntff-sizez", "5"],
    ["--same-size-topology"],                                                      # listed by --help, rejected by the parser
    ["--no-such-option"],
    ["--sizex", "40", "--bogus", "--sizey", "50"],
    ["--version"],
]


def main():
    out = []
    for argv in CASES:
        r = subprocess.run([EXE] + argv, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        out.append({"argv": argv, "rc": r.returncode, "stdout": r.stdout})
    # --cmd-from-file: the file holds one token per line
    with tempfile.TemporaryDirectory() as tmp:
        f = os.path.join(tmp, "cmd.txt")
        open(f, "w").write("--sizex\n48\n--same-size\n--use-pml\n--dx\n0.00025\n")
        r = subprocess.run([EXE, "--cmd-from-file", f], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        out.append({"argv": ["--cmd-from-file", "@FILE"], "file": "--sizex\n48\n--same-size\n--use-pml\n--dx\n0.00025\n",
                    "rc": r.returncode, "stdout": r.stdout.replace(f, "@FILE")})
        r = subprocess.run([EXE, "--cmd-from-file", f, "--use-tfsf"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        out.append({"argv": ["--cmd-from-file", "@FILE", "--use-tfsf"], "file": "--sizex\n48\n", "rc": r.returncode,
                    "stdout": r.stdout.replace(f, "@FILE")})
    json.dump(out, open(os.path.join(HERE, "settings_cases.json"), "w"), indent=1)
    for c in out:
        print(c["rc"], c["argv"])


if __name__ == "__main__":
    main()
