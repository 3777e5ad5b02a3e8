# chiaswarm_b200 on Windows: the sm_100a kernels are built with nvcc for Linux, so the worker runs inside WSL2
# (Ubuntu) with the NVIDIA driver's WSL CUDA support; this script checks the prerequisites and runs install.sh
# there.  Reference: Install.ps1:1-103.
$ErrorActionPreference = "Stop"

if (-not [Environment]::Is64BitOperatingSystem) {
    Write-Error "chiaswarm_b200 requires 64-bit Windows"
    exit 1
}
$wsl = Get-Command wsl.exe -ErrorAction SilentlyContinue
if (-not $wsl) {
    Write-Error "WSL2 is required: run 'wsl --install -d Ubuntu-24.04' from an elevated prompt, reboot, then re-run this script"
    exit 1
}
$distros = (wsl.exe -l -q) -replace "`0", "" | Where-Object { $_ -ne "" }
if (-not $distros) {
    Write-Error "no WSL distribution installed: wsl --install -d Ubuntu-24.04"
    exit 1
}
$distro = if ($env:SWARM_WSL_DISTRO) { $env:SWARM_WSL_DISTRO } else { $distros[0] }
Write-Output "using WSL distribution: $distro"

$smi = wsl.exe -d $distro -- bash -lc "command -v nvidia-smi >/dev/null && nvidia-smi --query-gpu=name --format=csv,noheader"
if ($LASTEXITCODE -ne 0) {
    Write-Error "the GPU is not visible inside WSL: install the current NVIDIA Windows driver (it carries the WSL CUDA libraries)"
    exit 1
}
Write-Output "GPU(s) inside WSL: $smi"

$here = (Get-Location).Path
$wslPath = (wsl.exe -d $distro -- wslpath -a "$here").Trim()
wsl.exe -d $distro -- bash -lc "cd '$wslPath' && ./install.sh"
if ($LASTEXITCODE -ne 0) {
    Write-Error "install.sh failed inside WSL"
    exit 1
}
Write-Output ""
Write-Output "Installed.  Start the worker with:"
Write-Output "  wsl -d $distro -- bash -lc `"cd '$wslPath' && . ./activate && python -m swarm.worker`""
