@ECHO OFF
REM Windows front end of the Sphinx build (same targets as the Makefile): make.bat html
pushd %~dp0
if "%SPHINXBUILD%" == "" set SPHINXBUILD=sphinx-build
set SOURCEDIR=.
set BUILDDIR=_build
%SPHINXBUILD% >NUL 2>NUL
if errorlevel 9009 (
	echo.sphinx-build was not found: pip install -r requirements.txt, or set SPHINXBUILD to its full path.
	exit /b 1
)
if "%1" == "" (
	%SPHINXBUILD% -M help %SOURCEDIR% %BUILDDIR% %SPHINXOPTS% %O%
) else (
	%SPHINXBUILD% -M %1 %SOURCEDIR% %BUILDDIR% %SPHINXOPTS% %O%
)
popd
