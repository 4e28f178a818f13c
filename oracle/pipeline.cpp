/* placeholder; filled below */
