//! A small stand-in for src/logger.rs so the driver's `Logger::new(path)` / `set_level` / `write`
//! lines keep compiling: "<unix seconds> [LEVEL] - message" appended to the file.  (Real progress
//! reporting is rtb_render_progressive's callback.)
use std::fs::OpenOptions;
use std::io::Write;
use std::time::{SystemTime, UNIX_EPOCH};

#[derive(Clone, Copy, Debug)]
#[allow(clippy::upper_case_acronyms)]
pub enum LogLevel {
    INFO,
    DEBUG,
    WARNING,
    ERROR,
}

#[derive(Clone, Debug)]
pub struct Logger {
    level: LogLevel,
    path: String,
}

impl Logger {
    pub fn new(path: &str) -> Self {
        Logger { level: LogLevel::DEBUG, path: path.to_owned() }
    }
    pub fn set_level(&mut self, l: LogLevel) {
        self.level = l
    }
    pub fn write(&self, description: &str) -> std::io::Result<()> {
        let secs = SystemTime::now().duration_since(UNIX_EPOCH).map(|d| d.as_secs()).unwrap_or(0);
        let mut f = OpenOptions::new().append(true).create(true).open(&self.path)?;
        writeln!(f, "{} [{:?}] - {}", secs, self.level, description)
    }
}
