module github.com/b200crc/b200crc

go 1.21
