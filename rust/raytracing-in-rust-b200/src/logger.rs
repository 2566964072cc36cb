//! Stand-in for the reference's logger module so the driver's `Logger::new(path)`, `set_level(DEBUG)` and
//! `write("..")` lines keep compiling.  One line per call is appended to the file:
//! "<unix seconds> [LEVEL] - message".  (Real progress reporting is rtb_render_progressive's callback.)
use std::fmt;
use std::io::Write as _;
use std::path::PathBuf;

#[allow(clippy::upper_case_acronyms)]
#[derive(Debug, Clone, Copy)]
pub enum LogLevel { INFO, DEBUG, WARNING, ERROR }

impl fmt::Display for LogLevel {
    fn fmt(&self, f: &mut fmt::Formatter<'_>) -> fmt::Result {
        fmt::Debug::fmt(self, f)
    }
}

#[derive(Debug, Clone)]
pub struct Logger { file: PathBuf, tag: LogLevel }

impl Logger {
    pub fn new(path: &str) -> Self {
        Logger { file: PathBuf::from(path), tag: LogLevel::DEBUG }
    }

    pub fn set_level(&mut self, l: LogLevel) {
        self.tag = l;
    }

    pub fn write(&self, description: &str) -> std::io::Result<()> {
        let now = std::time::SystemTime::now().duration_since(std::time::UNIX_EPOCH).map_or(0, |d| d.as_secs());
        let mut out = std::fs::OpenOptions::new().create(true).append(true).open(&self.file)?;
        writeln!(out, "{} [{}] - {}", now, self.tag, description)
    }
}
